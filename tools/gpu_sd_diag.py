"""Diagnostic (GPU box): vspg_debug_sample_distance / vspg_debug_guided_sample vs the oracle, field by field, bitwise."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
from pyvspg import api
import helpers as H
import orc
import test_gpu_parity as T
ctx = api.Context(0)
f = lambda a, k: np.asarray(H.struct_array_to_np(a, k))
def b32(x): return np.ascontiguousarray(np.asarray(x, dtype=np.float32)).view(np.uint32)
for name, over in T.SD_CASES:
    cfg = H.small(name)
    for k, v in over.items(): setattr(cfg.ip, k, v)
    T._setup(ctx, cfg)
    rng = np.random.default_rng(17); n = 3000
    inp = T._sd_inputs(cfg, rng, n)
    got = ctx.debug_sample_distance(inp); want = orc.sample_distance(cfg.scene, cfg.guiding, cfg.ip, inp)
    msg = []
    for k in ("scattered", "terminated", "depth", "rng_state", "density_queries", "dda_steps"):
        msg.append(f"{k}:{int((f(got, k) != f(want, k)).sum())}")
    for k in ("p", "sigma_a", "sigma_s", "beta", "r_u", "r_l"):
        a, b = b32(f(got, k)), b32(f(want, k))
        nanboth = np.isnan(f(got, k)) & np.isnan(f(want, k))
        bad = ((a != b) & ~nanboth).any(axis=1)
        msg.append(f"{k}:{int(bad.sum())}")
        if bad.any() and k in ("r_u", "beta", "p"):
            i = np.where(bad)[0][0]
            msg.append(f"[{i}: {f(got, k)[i]} vs {f(want, k)[i]}]")
    print(name, over, " ".join(msg))
