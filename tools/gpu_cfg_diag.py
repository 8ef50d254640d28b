"""Diagnostic (GPU box): CUDA vs oracle on a pyvspg config, per (pixel, sample), with integrator overrides."""
import os, sys, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
from pyvspg import api
import helpers as H
import orc
ctx = api.Context(0)
PF = C.POINTER(C.c_float)
name = sys.argv[1]
VARIANTS = [("full", {}), ("nonee", dict(use_nee=0)), ("nobias", dict(collision_probability_bias=0)), ("resampling", dict(vsp_sampling_nds=0, collision_probability_bias=0)),
            ("noguide", dict(surface_guiding=0, volume_guiding=0)), ("novsp", dict(vsp_guiding=0)), ("d1", dict(max_depth=1))]
for vn, over in VARIANTS:
    cfg = H.small(name)
    for k, v in over.items(): setattr(cfg.ip, k, v)
    ctx.set_scene(cfg.scene); ctx.set_guiding(cfg.guiding); ctx.set_integrator(cfg.ip)
    W, Hh = cfg.width, cfg.height
    gp = C.byref(cfg.guiding.desc) if cfg.guiding is not None else None
    nbad = 0; tot = 0; shown = 0
    for s in range(cfg.spp):
        ctx.film_clear(); ctx.render(s, s + 1, seed=7); film = ctx.film_download()[..., :3].astype(np.float32)
        fo, _ = orc.render(cfg.scene, cfg.guiding, cfg.ip, s, s + 1, seed=7, nthreads=16)
        fo = fo[..., :3].astype(np.float32)
        bad = (film.view(np.uint32) != fo.view(np.uint32)).any(axis=2)
        nbad += bad.sum(); tot += bad.size
        for (y, x) in np.argwhere(bad)[:2]:
            if shown < 4: print("     ", vn, (x, y, s), film[y, x], fo[y, x]); shown += 1
    print(f"{name} {vn:11s}: {nbad} of {tot} samples differ")
