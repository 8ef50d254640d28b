"""Small renders of every kernel family (written for compute-sanitizer memcheck, which is closed on this pool; kept as a quick
all-kernels smoke run): guided grid (C2), NDS+ (C3), rgbgrid with all
light types (C5), homogeneous (C1), NanoVDB, triangles / spheres; capacity below the sample count so that regeneration
and band-major ordering run; the queue coherence pass on; the allocation-free hook."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("VSPG_SORT", "3")
import numpy as np
import helpers as H
from pyvspg import api
ctx = api.Context(0)
for name in ["C1", "C1T", "C1S", "C1E", "C2", "C3", "C4", "C5"]:
    cfg = H.small(name, res=24, spp=2)
    ctx.set_scene(cfg.scene); ctx.set_guiding(cfg.guiding); ctx.set_integrator(cfg.ip)
    ctx.set_capacity(1024)
    ctx.film_clear(); ctx.render(0, cfg.spp, seed=1)
    f = ctx.film_download()
    rays = H.random_rays(np.random.default_rng(1), 256, cfg.scene.eye)
    if cfg.scene.n_media:
        ctx.debug_majorant_segments_noalloc(0, rays, 32)
    print(name, "ok", float(f[..., :3].sum()))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_nanovdb as T
sc = T.nvdb_scene()
ctx.set_scene(sc); ctx.set_guiding(None); ctx.set_integrator(api.integrator_defaults(vsp_guiding=0, surface_guiding=0, volume_guiding=0))
ctx.film_clear(); ctx.render(0, 2, seed=1)
print("nanovdb ok", float(ctx.film_download()[..., :3].sum()))
ctx.close()
