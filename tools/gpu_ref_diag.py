"""Diagnostic (GPU box): per-scene agreement of the CUDA path with the reference's golden per-sample radiance."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
from pyvspg import api
from ref_scenes import SCENES
GOLD = os.path.join(ROOT, "tests", "golden", "pbrt")
ctx = api.Context(0)
names = sys.argv[1:] or SCENES
for name in names:
    hs = api.HostScene(path=os.path.join(GOLD, name + ".pbrt")); g = np.load(os.path.join(GOLD, name + ".npz"))
    variants = [("full", {})]
    if "--bisect" in os.environ.get("DIAG", ""):
        variants += [("nonee", dict(use_nee=0))]
    for vn, over in variants:
        ctx.set_scene(hs)
        if hs.guiding is not None: ctx.set_guiding(hs.guiding)
        ip = hs.ip
        ctx.set_integrator(ip)
        smp, Lref, w = g["samples"], g["L"], g["rays"][:, 7]
        films = {}
        for s in range(hs.spp):
            ctx.film_clear(); ctx.render(s, s + 1, seed=hs.seed); films[s] = ctx.film_download().copy()
        got = np.stack([films[int(s)][int(py), int(px), :3] for px, py, s in smp]).astype(np.float32)
        want = (w[:, None] * Lref).astype(np.float32)
        eq = (got.view(np.uint32) == want.view(np.uint32)).all(axis=1)
        close = np.isclose(got, want, rtol=1e-5, atol=1e-8).all(axis=1)
        print(f"{name:10s} {vn}: bit-equal {eq.mean():.4f}  within1e-5 {close.mean():.4f}  mean got {got.mean():.6f} want {want.mean():.6f}")
        bad = np.where(~eq)[0][:3]
        for i in bad: print("     ", smp[i], got[i], want[i])
    hs.close()
