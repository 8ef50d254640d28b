"""Debug aid: per-case mismatch report of vspg_debug_sample_distance vs the oracle (run on the GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("vspg-pbrt-v4_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import numpy as np
import helpers as H, orc
from pyvspg import api
import test_gpu_parity as T

ctx = api.Context(0)
f = lambda a, k: H.struct_array_to_np(a, k)
for name, over in T.SD_CASES:
    cfg = H.small(name)
    for k, v in over.items():
        setattr(cfg.ip, k, v)
    T._setup(ctx, cfg)
    rng = np.random.default_rng(17)
    n = 3000
    inp = T._sd_inputs(cfg, rng, n)
    got = ctx.debug_sample_distance(inp)
    want = orc.sample_distance(cfg.scene, cfg.guiding, cfg.ip, inp)
    keys = ("scattered", "terminated", "depth", "rng_state", "density_queries", "dda_steps")
    same = np.ones(n, bool)
    msg = []
    for k in keys:
        eq = f(got, k) == f(want, k)
        msg.append(f"{k}:{(~eq).sum()}")
        same &= eq
    print(f"== {name} {over}: decisions mismatch {(~same).sum()}/{n}  [{' '.join(msg)}] scattered={int(f(want,'scattered').sum())}")
    for k in ("p", "sigma_a", "sigma_s", "beta", "r_u", "r_l"):
        a, b = f(got, k)[same], f(want, k)[same]
        fin = np.isfinite(b).all(axis=1)
        finm = (np.isfinite(a).all(axis=1) == fin).all()
        if k == "p":
            err = np.abs(a[fin] - b[fin]) / np.maximum(np.abs(b[fin]).max(axis=1, keepdims=True), 1e-12)
        else:
            err = H.rel_err(a[fin], b[fin], floor=1e-12)
        w = np.unravel_index(np.argmax(err), err.shape)
        print(f"   {k:8s} finite-match={finm} max={err.max():.3e} n>1e-5={(err.max(axis=1) > 1e-5).sum()}  worst got={a[fin][w[0]]} want={b[fin][w[0]]}")
    bad = np.where(~same)[0][:3]
    for i in bad:
        I = inp[i]
        print(f"   bad[{i}] depth={I.depth} ch={I.channel} vsp={I.vsp:.4f} tmax={I.t_max:.4f} got(sc={got[i].scattered},term={got[i].terminated},q={got[i].density_queries},dda={got[i].dda_steps}) "
              f"want(sc={want[i].scattered},term={want[i].terminated},q={want[i].density_queries},dda={want[i].dda_steps})")
