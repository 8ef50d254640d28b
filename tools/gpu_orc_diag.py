"""Diagnostic (GPU box): CUDA path vs the CPU oracle (itself bit-equal to the reference), per (pixel, sample), under
integrator-parameter overrides -- to bisect which stage introduces last-bit differences."""
import os, sys, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
from pyvspg import api
import orc
GOLD = os.path.join(ROOT, "tests", "golden", "pbrt")
ctx = api.Context(0)
PF = C.POINTER(C.c_float)
VARIANTS = [("full", {}), ("nonee", dict(use_nee=0)), ("nee_d1", dict(max_depth=1)), ("nee_d2", dict(max_depth=2)),
            ("noguide", dict(surface_guiding=0, volume_guiding=0)), ("novsp", dict(vsp_guiding=0)), ("uniform", dict(light_sampler=0)), ("power", dict(light_sampler=1)),
            ("noguide_novsp", dict(surface_guiding=0, volume_guiding=0, vsp_guiding=0))]
for name in sys.argv[1:]:
    hs = api.HostScene(path=os.path.join(GOLD, name + ".pbrt")); g = np.load(os.path.join(GOLD, name + ".npz"))
    ctx.set_scene(hs)
    if hs.guiding is not None: ctx.set_guiding(hs.guiding)
    smp, w = g["samples"], g["rays"][:, 7]
    gp = C.byref(hs.guiding.desc) if hs.guiding is not None else None
    for vn, over in VARIANTS:
        ip = api.integrator_defaults()
        C.memmove(C.byref(ip), C.byref(hs.ip), C.sizeof(ip))
        for k, v in over.items(): setattr(ip, k, v)
        ctx.set_integrator(ip)
        films = {}
        for s in range(hs.spp):
            ctx.film_clear(); ctx.render(s, s + 1, seed=hs.seed); films[s] = ctx.film_download().copy()
        got = np.stack([films[int(s)][int(py), int(px), :3] for px, py, s in smp]).astype(np.float32)
        want = np.zeros((len(smp), 3), np.float32); o = np.zeros(3, np.float32)
        for i, (px, py, s) in enumerate(smp):
            orc.lib().orc_render_pixel_sample(C.byref(hs.desc), gp, C.byref(ip), int(px), int(py), int(s), hs.seed, o.ctypes.data_as(PF)); want[i] = w[i] * o
        eq = (got.view(np.uint32) == want.view(np.uint32)).all(axis=1)
        close = np.isclose(got, want, rtol=1e-5, atol=1e-8).all(axis=1)
        print(f"{name:8s} {vn:9s}: bit-equal {eq.mean():.4f}  within1e-5 {close.mean():.4f}  nonblack {(want != 0).any(axis=1).mean():.3f}")
        for i in np.where(~eq)[0][:2]: print("     ", smp[i], got[i], want[i])
    hs.close()
