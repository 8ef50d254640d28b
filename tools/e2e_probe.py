import sys, time, os
ROOT = os.environ.get("GRAFT_REPO_ROOT", "/root/repo")
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200"))
import torch
from pyvspg import api, configs
cfg = configs.BUILDERS["C2"]()
ctx = api.Context()
W, H = cfg.width, cfg.height
film_host = torch.empty((H, W, 4), dtype=torch.float64).pin_memory()
for it in range(8):
    t = [time.perf_counter()]
    ctx.set_scene(cfg.scene); torch.cuda.synchronize(); t.append(time.perf_counter())
    ctx.set_guiding(cfg.guiding); torch.cuda.synchronize(); t.append(time.perf_counter())
    ctx.set_integrator(cfg.ip); t.append(time.perf_counter())
    ctx.film_clear(); ctx.render(0, cfg.spp, seed=it); torch.cuda.synchronize(); t.append(time.perf_counter())
    ctx.film_download_to(film_host.data_ptr()); t.append(time.perf_counter())
    print(it, " ".join(f"{(b - a) * 1e3:8.2f}" for a, b in zip(t, t[1:])), flush=True)
