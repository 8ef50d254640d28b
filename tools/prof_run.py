"""Lean profiling driver: N renders of a workload through the C ABI (no CPU legs, no e2e), for ncu."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200"))
from pyvspg import api, configs
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2
kw = {}
if len(sys.argv) > 3: kw["spp"] = int(sys.argv[3])
cfg = configs.BUILDERS[name](**kw)
ctx = api.Context(0)
ctx.set_scene(cfg.scene); ctx.set_guiding(cfg.guiding); ctx.set_integrator(cfg.ip)
for i in range(n):
    ctx.film_clear(); ctx.render(0, cfg.spp, seed=i)
f = ctx.film_download()
print("ok", name, n, float(f[..., :3].sum() / max(1.0, f[..., 3].sum())), ctx.counters()["kernel_launches"])
ctx.close()
