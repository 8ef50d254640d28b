"""Debug aid: per-pixel GPU vs oracle film comparison under parameter sweeps (run on the GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("vspg-pbrt-v4_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import numpy as np
import helpers as H, orc
from pyvspg import api

ctx = api.Context(0)
names = sys.argv[1:] or ["C1", "C2"]
for name in names:
    for over in (dict(max_depth=2), dict(max_depth=2, min_rr_depth=10), dict(max_depth=3, min_rr_depth=10), dict(min_rr_depth=10)):
        cfg = H.small(name, spp=1)
        for k, v in over.items():
            setattr(cfg.ip, k, v)
        ctx.set_scene(cfg.scene); ctx.set_guiding(cfg.guiding); ctx.set_integrator(cfg.ip)
        ctx.film_clear(); ctx.counters_reset()
        ctx.render(0, 1, seed=7)
        a = ctx.film_download(); cg = ctx.counters()
        b, co = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, 1, seed=7, nthreads=8)
        err = H.rel_err(a[..., :3], b[..., :3], floor=1e-3).max(axis=2)
        bad = np.argwhere(err > 1e-3)
        print(f"== {name} {over}: wsum_equal={np.array_equal(a[...,3], b[...,3])} bad={len(bad)}/{err.size} mean g={a[...,:3].mean():.6f} o={b[...,:3].mean():.6f}")
        print("   counters g:", {k: cg[k] for k in ('density_queries','dda_steps','cache_lookups','shadow_rays','bvh_nodes')})
        print("   counters o:", co)
        for (y, x) in bad[:6]:
            print(f"   px({x},{y}) g={a[y,x,:3]} o={b[y,x,:3]}")
