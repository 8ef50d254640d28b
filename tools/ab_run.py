"""A/B timing driver: renders a workload N times through the C ABI and prints ms per render (CUDA events inside the
library) plus the per-stage split, and a checksum of the film.  The film is summed with fp64 atomics in completion order, so
the checksum of a full-size render can differ between two runs of the SAME build in a few last-bit roundings (C3: a handful of
distinct values); identity of variants is what the pixel tests against the oracle establish, not this hash.
    VSPG_LIB=/path/to/libvspg_variant.so python tools/ab_run.py C2 3"""
import hashlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200"))
import numpy as np
from pyvspg import api, configs
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
kw = {}
if len(sys.argv) > 3: kw["spp"] = int(sys.argv[3])
cfg = configs.BUILDERS[name](**kw)
ctx = api.Context(0)
if os.environ.get("VSPG_AB_CAPACITY"): ctx.set_capacity(int(os.environ["VSPG_AB_CAPACITY"]))
ctx.set_scene(cfg.scene); ctx.set_guiding(cfg.guiding); ctx.set_integrator(cfg.ip)
ctx.film_clear(); ctx.render(0, cfg.spp, seed=0)           # warm-up
ctx.kernel_stats(True)                                     # switch stage timing on
ctx.counters_reset()
for i in range(n):
    ctx.film_clear(); ctx.render(0, cfg.spp, seed=0)
st = ctx.kernel_stats(True)
f = ctx.film_download()
h = hashlib.sha1(np.ascontiguousarray(f).tobytes()).hexdigest()[:12]
tot = st["ms_total"]
print("AB", os.path.basename(os.environ.get("VSPG_LIB", "libvspg.so")), name, "ms/render %.2f" % (tot / n), "film", h,
      " ".join("%s=%.2f" % (k, v["ms"] / n) for k, v in st.items() if isinstance(v, dict) and v["ms"] > 0))
ctx.close()
