#!/usr/bin/env python
"""bench.py -- paths/sec of the guidedvolpath(VSPG) render pass on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C2] [--impl reference]

A *step* is one render of the whole workload (every (pixel, sampleIndex) of the configuration) on each rank.
Workload = BASELINE.json configs[1] (C2: fBm 128^3 grid medium, VSPG resampling, loaded cache + ISG buffer,
512x512, 64 spp) unless --workload says otherwise.  N > 1 (launched by torchrun, one rank per GPU): sample-pass
sharding -- rank r renders sample indices r, r+N, ... of an N*spp render (per-GPU work fixed => "weak"), and the fp64
film buffers are combined with one NCCL reduce over NVLink inside the timed region.

JSON line keys (see DESIGN.md "Measurement"):
  value     whole-job paths/s, inputs already resident in HBM (CUDA events, max over ranks)
  e2e       same metric through the C ABI from HOST buffers: set_scene + set_guiding (H2D of grids / cache / buffers),
            render, film download (D2H) inside the timed region
  roofline  the dominant kernel's algorithmic grid bytes (Q*b_q + S*4) / its CUDA-event time vs the
            measured HBM copy peak (MEASURED_PEAKS.json)
  cpu_baseline  the reference renderer (oracle/_ref, kind "reference") on this box's host cores, bounded sample
--impl reference: the REFERENCE RENDERER ITSELF (oracle/_ref: its own sources, compiled by oracle/build_ref_pbrt.sh) runs
the workload from the scene / cache / buffer files on all host cores, same metric/config; nothing of libvspg.so is loaded.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200"))

WORKLOADS = {
    # name: (builder key, kwargs, b_q bytes per density query)
    "C1": ("C1", {}, 0),
    "C2": ("C2", {}, 32),
    "C3": ("C3", {}, 32),
    "C4": ("C4", {}, 32),
    "C5": ("C5", {}, 192),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (reported in config)")
    ap.add_argument("--res-scale", type=float, default=1.0, help="scale film resolution (reported in config)")
    ap.add_argument("--capacity", type=int, default=0, help="paths in flight (0 = library default)")
    ap.add_argument("--shard", default="samples", choices=["samples", "bands"], help="multi-GPU partition (pyvspg.multi)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank renders the workload's spp (job = spp x N); strong: the workload's spp is split over the ranks")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-spp", type=int, default=0, help="spp of the CPU sample (0 = auto, ~10-30 s)")
    return ap.parse_args()


def build_config(args):
    from pyvspg import configs
    key, kw, bq = WORKLOADS[args.workload]
    kw = dict(kw)
    b = configs.BUILDERS[key]
    if args.spp:
        kw["spp"] = args.spp
    if args.res_scale != 1.0:
        import inspect
        sig = inspect.signature(b).parameters
        if "res" in sig:
            kw["res"] = int(sig["res"].default * args.res_scale)
        else:
            kw["width"] = int(sig["width"].default * args.res_scale)
            kw["height"] = int(sig["height"].default * args.res_scale)
    return b(**kw), bq


def config_dict(cfg, args, world):
    grid = "homogeneous"
    for kind, e in cfg.scene.log:
        if kind == "medium" and e["type"] != "homogeneous":
            g = e["density"] if e["type"] == "uniformgrid" else e["sigma_a_rgb"]
            grid = [int(g.shape[2]), int(g.shape[1]), int(g.shape[0])]
            break
    return {
        "workload": cfg.name, "description": cfg.description, "width": cfg.width, "height": cfg.height,
        "spp_per_gpu": cfg.spp if args.scaling == "weak" else cfg.spp / world, "spp_total": cfg.spp * (world if args.scaling == "weak" else 1), "grid": grid,
        "sampler": "independent", "maxdepth": cfg.ip.max_depth,
        "parallelism": f"{'sample-pass' if args.shard == 'samples' else 'interleaved-band'} sharding x{world}" + (" + NCCL film reduce" if world > 1 else ""),
        "l2": "inputs larger than L2: the path-state working set (up to 2^24 paths x ~380 B = 6.4 GB) streams through the 126 MB L2 every wave "
              "iteration; the density cell records are 67 MB for 128^3 (L2-sized, evicted by the state traffic) and 4.3 GB for 512^3",
    }


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs (B200_PROFILING.md)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop = [], set(), None, threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # NVML missing: record that, do not fail the bench
            self.nv = None
            self.err = str(e)

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "NVML unavailable: " + getattr(self, "err", "")}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


REF_DRIVER = r"""
import ctypes as C, json, os, sys
L = C.CDLL(sys.argv[1])                      # oracle/_ref/libpbrtharness.so (+ libpbrt_ref.so: the reference's own code)
L.refp_open.restype = C.c_void_p; L.refp_open.argtypes = [C.c_char_p]
L.refp_render.restype = C.c_double; L.refp_render.argtypes = [C.c_void_p, C.c_void_p]
L.refp_init(int(sys.argv[3]), 0)             # --nthreads N, rendering space "cameraworld" (the reference's default)
h = L.refp_open(sys.argv[2].encode())
t = L.refp_render(h, None)                   # integrator->Render(), timed around the call
sys.stdout.flush()
os.write(1, ("\nREFP_RESULT " + json.dumps({"render_s": t}) + "\n").encode())
os._exit(0)
"""


def reference_available():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libpbrtharness.so")) and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libpbrt_ref.so"))


class ReferenceArm:
    """The REFERENCE RENDERER ITSELF on this box's host cores: oracle/_ref/libpbrt_ref.so is the reference's own source
    compiled by oracle/build_ref_pbrt.sh; this class writes the workload as the files the reference's scene language
    names (scene.pbrt, guiding cache, ISG / Tr buffers) and runs GuidedVolPathVSPGIntegrator::Render() on them in a
    child process with all host threads.  Nothing of the product (libvspg.so, CUDA) is loaded on this path."""

    def __init__(self, cfg):
        import tempfile
        from pyvspg import export
        self.cfg = cfg
        self.dir = tempfile.mkdtemp(prefix="vspg_ref_")
        self.export = export
        self.cores = os.cpu_count() or 1
        self.template = os.path.join(ROOT, "tests", "golden", "field_small.pgl")
        self._written = {}

    def scene(self, spp):
        if spp not in self._written:
            self._written[spp] = self.export.export_config(self.cfg, self.dir, self.template, name=f"w{spp}", outfile=f"w{spp}.pfm", spp=spp)
        return self._written[spp]

    def render(self, spp):
        """One Render() of W x H x spp paths; returns (paths, seconds inside Render())."""
        import subprocess
        path = self.scene(spp)
        r = subprocess.run([sys.executable, "-c", REF_DRIVER, os.path.join(ROOT, "oracle", "_ref", "libpbrtharness.so"), path, str(self.cores)],
                           cwd=self.dir, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        for line in r.stdout.splitlines():
            if line.startswith("REFP_RESULT "):
                return self.cfg.width * self.cfg.height * spp, json.loads(line[len("REFP_RESULT "):])["render_s"]
        raise RuntimeError("reference render failed:\n" + r.stdout[-2000:])

    def close(self):
        import shutil
        shutil.rmtree(self.dir, ignore_errors=True)


def cpu_oracle_paths_per_s(cfg, spp, seed=0):
    """Fallback when oracle/_ref is absent: the CPU oracle port (oracle/liboracle.so) on all host cores.
    Checker/baseline use only: nothing on the product path calls this."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    orc.lib()
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    _, cnt = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, spp, seed=seed, nthreads=cores)
    dt = time.perf_counter() - t0
    return cnt["paths"] / dt, cores, dt, cnt["paths"]


def cpu_reference_sample(cfg, target_s, fixed_spp=0):
    """A bounded sample of the workload on the host cores: (value paths/s, cores, kind, sample text)."""
    if reference_available() and os.path.exists(os.path.join(ROOT, "tests", "golden", "field_small.pgl")):
        arm = ReferenceArm(cfg)
        try:
            n1, t1 = arm.render(1)                                   # calibration (and page-in) pass
            spp = fixed_spp or max(1, min(cfg.spp, int(target_s / max(t1, 1e-3))))
            n, t = arm.render(spp)
        finally:
            arm.close()
        return n / t, arm.cores, "reference", (f"{cfg.width}x{cfg.height} x {spp} spp of the {cfg.spp}-spp workload ({n} paths, {t:.1f} s inside "
                                               f"GuidedVolPathVSPGIntegrator::Render()); the reference's own sources built by oracle/build_ref_pbrt.sh, --nthreads {arm.cores}")
    v1, cores, dt1, _ = cpu_oracle_paths_per_s(cfg, 1)
    spp = fixed_spp or max(1, min(cfg.spp, int(target_s / max(dt1, 1e-3))))
    v, cores, dt, n = cpu_oracle_paths_per_s(cfg, spp, seed=1)
    return v, cores, "port", (f"{cfg.width}x{cfg.height} x {spp} spp of the {cfg.spp}-spp workload ({n} paths, {dt:.1f} s); oracle/liboracle.so "
                              "(CPU restatement; oracle/_ref not present), all host threads")


def relmse_check(ctx, args):
    """SURVEY 8(d) "relMSE vs CPU ref at equal spp", on a bounded instance of the workload (same scene, 128-pixel film):
    relMSE of the CUDA render and of the CPU oracle render against a common high-spp reference rendered BY THE CPU ORACLE
    (so a bias of the CUDA path cannot hide in the reference), 8 seeds each.  The oracle is the checker here (cpu leg of
    the bench only); it is itself pinned bit for bit to the reference renderer (tests/test_ref_pbrt.py)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    from pyvspg import configs, stats
    key, kw, _ = WORKLOADS[args.workload]
    small = dict(C1=dict(res=128, spp=8), C2=dict(res=128, spp=8, grid_n=64, vol_leaves=512, surf_leaves=256),
                 C3=dict(res=128, spp=8, grid_n=64, vol_leaves=512, surf_leaves=256),
                 C4=dict(width=160, height=90, spp=8, grid_n=64, vol_leaves=512, surf_leaves=256),
                 C5=dict(width=160, height=90, spp=8, grid_n=32, vol_leaves=512, surf_leaves=256))[key]
    cfg = configs.BUILDERS[key](**small)
    cores = os.cpu_count() or 1
    ref_spp = 512
    film_ref, _ = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, ref_spp, seed=4242, nthreads=cores)
    ref = stats.resolve(film_ref)
    ctx.set_scene(cfg.scene); ctx.set_guiding(cfg.guiding); ctx.set_integrator(cfg.ip)
    rg, rc = [], []
    for seed in range(8):
        ctx.film_clear(); ctx.render(0, cfg.spp, seed=seed)
        rg.append(stats.rel_mse(stats.resolve(ctx.film_download()), ref))
        film_o, _ = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, cfg.spp, seed=100 + seed, nthreads=cores)
        rc.append(stats.rel_mse(stats.resolve(film_o), ref))
    r = stats.compare(rg, rc)
    return {"gpu": r["a"], "cpu": r["b"], "abs_diff": r["diff"], "tolerance": r["tolerance"], "agree": r["agree"], "seeds": 8,
            "instance": f"{cfg.name} at {cfg.width}x{cfg.height}, {cfg.spp} spp; reference = {ref_spp} spp rendered by the CPU oracle (disjoint seed)"}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores (oracle/_ref when it was
    built -- kind "reference" --, else the oracle port), same metric / config; each step is a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg, _ = build_config(args)
    cores = os.cpu_count() or 1
    if reference_available():
        arm = ReferenceArm(cfg)
        try:
            n1, t1 = arm.render(1)
            spp_step = args.cpu_spp or max(1, min(cfg.spp, int(8.0 / max(t1, 1e-3))))
            for _ in range(max(0, min(args.warmup, 3) - 1)):
                arm.render(1)
            tot_paths, tot_t = 0, 0.0
            for s in range(args.steps):
                n, t = arm.render(spp_step)
                tot_paths += n
                tot_t += t
        finally:
            arm.close()
        kind = "reference"
        how = "the reference's own sources (oracle/_ref, built by oracle/build_ref_pbrt.sh), Render() timed, --nthreads %d" % cores
    else:
        v1, cores, dt1, _ = cpu_oracle_paths_per_s(cfg, 1)
        spp_step = args.cpu_spp or max(1, min(cfg.spp, int(8.0 / max(dt1, 1e-3))))
        tot_paths, tot_t = 0, 0.0
        for s in range(args.steps):
            v, cores, dt, n = cpu_oracle_paths_per_s(cfg, spp_step, seed=s)
            tot_paths += n
            tot_t += dt
        kind = "port"
        how = "oracle/liboracle.so (CPU restatement; oracle/_ref not present)"
    val = tot_paths / tot_t
    sample = f"{cfg.width}x{cfg.height} x {spp_step} spp per step of the {cfg.spp}-spp workload (sample passes are independent); {how}"
    print(json.dumps({
        "impl": "reference", "metric": "paths/sec (guidedvolpathvspg render pass)", "value": val, "unit": "paths/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(1, args.steps), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(cfg, args, 1),
        "cpu_baseline": {"value": val, "unit": "paths/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "paths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


class _DevArray:
    """Zero-copy view of ctx-owned device memory for torch (NCCL reduce operates on the film buffer in place)."""

    def __init__(self, ptr, n, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def input_bytes(cfg):
    """Host->device bytes of one vspg_set_scene + vspg_set_guiding (what the e2e timed region copies)."""
    import ctypes as C
    from pyvspg import _abi as A
    n = 0
    d = cfg.scene.desc
    for i in range(d.n_media):
        m = d.media[i]
        nv = m.nx * m.ny * m.nz
        if m.type == A.MEDIUM_GRID:
            n += nv * 4 + 16 ** 3 * 4
        elif m.type == A.MEDIUM_RGBGRID:
            n += nv * 24 + 16 ** 3 * 4
    n += d.n_quads * C.sizeof(A.Quad) + d.n_lights * C.sizeof(A.Light) + d.n_materials * C.sizeof(A.Material)
    g = cfg.guiding
    if g is not None:
        for f in g.fields.values():
            n += f.n_regions * (C.sizeof(A.Mixture) + C.sizeof(A.Neighbours)) + f.n_nodes * C.sizeof(A.KdNode)
        if getattr(g, "isg", None) is not None:
            n += g.isg.nbytes
        if getattr(g, "tr", None) is not None:
            n += g.tr.nbytes
    return n


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from pyvspg import api, multi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the render pass has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the one JSON line ("NCCL version ..." goes to stderr)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cfg, b_q = build_config(args)
    ctx = api.Context(local)
    if args.capacity:
        ctx.set_capacity(args.capacity)
    ctx.set_scene(cfg.scene)
    ctx.set_guiding(cfg.guiding)
    ctx.set_integrator(cfg.ip)
    stream = torch.cuda.current_stream().cuda_stream
    fptr, fn = ctx.film_device_ptr()
    film_t = torch.as_tensor(_DevArray(fptr, fn), device=f"cuda:{local}")
    W, Hh, spp = cfg.width, cfg.height, cfg.spp
    spp_job = spp * world if args.scaling == "weak" else spp      # sample indices of one step, over all ranks
    paths_per_step = W * Hh * spp_job

    def render(b, e, s, window, seed):
        ctx.render(b, e, spp_stride=s, window=window, seed=seed, stream=stream)

    def step(seed):
        # weak scaling: a (spp*world)-spp render of the image; rank r takes its shard (sample indices r, r+world, ...
        # or its interleaved row bands), then the fp64 films are summed onto rank 0 (the one exchange step)
        ctx.film_clear()
        multi.render_shard(render, W, Hh, spp_job, rank, world, seed, mode=args.shard)
        multi.combine_film(film_t, world, dst=0)

    for i in range(args.warmup):
        step(i)
    barrier()
    ctx.kernel_stats(enable_timing=False)     # the headline is timed without the per-kernel events / syncs
    clocks = ClockSampler(local)
    clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for i in range(args.steps):
        step(100 + i)
    ev1.record()
    barrier()
    clk = clocks.stop()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=f"cuda:{local}", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    value = paths_per_step * args.steps / (ms_total * 1e-3)
    # per-kernel shares / roofline inputs: a separate pass of the same steps with the stage events on
    ctx.counters_reset()
    ctx.kernel_stats(enable_timing=True)
    barrier()
    sev0, sev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sev0.record()
    stat_steps = max(1, min(args.steps, 3))
    for i in range(stat_steps):
        step(100 + i)
    sev1.record()
    barrier()
    ms_stats = sev0.elapsed_time(sev1)
    cnt = ctx.counters()
    ks = ctx.kernel_stats(enable_timing=False)

    # ---- e2e: host buffers -> C ABI -> host film, copies inside the timed region
    h2d = input_bytes(cfg)
    film_host = torch.empty((Hh, W, 4), dtype=torch.float64).pin_memory()
    d2h = film_host.numel() * 8
    e2e_steps = max(1, min(args.steps, 3))

    def e2e_step(seed):
        ctx.set_scene(cfg.scene)          # H2D: density grid(s), majorant grid, shapes, lights
        ctx.set_guiding(cfg.guiding)      # H2D: guiding cache (KD nodes, mixtures, neighbour tables), ISG / Tr buffers
        ctx.set_integrator(cfg.ip)
        multi.render_shard(render, W, Hh, spp_job, rank, world, seed, mode=args.shard)
        multi.combine_film(film_t, world, dst=0)
        ctx.film_download_to(film_host.data_ptr())   # D2H: fp64 film

    for i in range(max(3, min(args.warmup, 5))):   # the first set_scene calls pay cudaMalloc of the staged blocks; later ones reuse them
        e2e_step(7 + i)
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step(200 + i)
    barrier()
    dt = torch.tensor([time.perf_counter() - t0], device=f"cuda:{local}", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = paths_per_step * e2e_steps / float(dt.item())

    if rank == 0:
        # ---- roofline of the dominant kernel (by summed CUDA-event time)
        names = [k for k in ks if k != "ms_total"]
        dom = max(names, key=lambda k: ks[k]["ms"])
        tracked = dom if ks[dom]["density_queries"] else max(("medium", "shadow"), key=lambda k: ks[k]["ms"])
        kk = ks[tracked]
        alg_bytes = kk["density_queries"] * b_q + kk["dda_steps"] * 4
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        achieved = alg_bytes / max(kk["ms"] * 1e-3, 1e-12) / 1e9 if kk["launches"] else 0.0
        per_launch_ms = kk["ms"] / max(1, kk["launches"])
        roofline = {
            "bound": "hbm", "kernel": "k_" + tracked, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": None, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": alg_bytes / max(1, kk["launches"]), "avg_launch_ms": per_launch_ms,
            "bytes_per_query": b_q, "queries": kk["density_queries"], "dda_steps": kk["dda_steps"],
            "kernel_ms_share": {k: round(ks[k]["ms"] / max(1e-9, sum(ks[n]["ms"] for n in names)), 4) for k in names},
            "kernel_ms_sum": sum(ks[n]["ms"] for n in names), "stats_pass_ms": ms_stats, "stats_pass_steps": stat_steps, "timed_region_ms": ms_total,
        }
        # ---- what only a profiler can count: DRAM bytes and issued instructions, from the committed ncu metric pass of this
        # workload (tools/ncu_metrics_summary.py -> profiles/); the durations they are divided by are this run's own
        prof_path = os.path.join(ROOT, "profiles", f"r02_ncu_metrics_{args.workload}.json")
        if os.path.exists(prof_path) and args.res_scale == 1.0:
            prof = json.load(open(prof_path))
            # the profiled render may have fewer spp than this run (ncu replays every launch): instruction counts scale
            # with the number of paths, per-launch DRAM bytes are quoted for the profiled wave size
            inst_scale = (cfg.spp / prof["spp"]) if prof.get("spp") else 1.0
            dk = max((k for k in prof["kernels"] if k.startswith("k_" + tracked)), key=lambda k: prof["kernels"][k]["ms"], default=None)
            if dk:
                roofline["traffic"] = prof["kernels"][dk]["dram_bytes_per_launch"]
                roofline["traffic_kernel"] = dk
                roofline["traffic_source"] = os.path.relpath(prof_path, ROOT) + (f" (profiled at {prof['spp']} spp)" if prof.get("spp") else "")
            big = max(names, key=lambda k: ks[k]["ms"])
            if big in prof["stages"]:
                ps_ = prof["stages"][big]
                sm_mhz = (clk or {}).get("sm_mhz") or 1965.0
                issue_peak = 148 * 4 * sm_mhz * 1e6          # warp instructions / s: 148 SMs x 4 schedulers x SM clock under load
                ach = ps_["warp_inst"] * inst_scale / max(1e-12, ks[big]["ms"] / stat_steps * 1e-3)
                roofline["issue"] = {"stage": big, "warp_inst_per_step": ps_["warp_inst"] * inst_scale, "lanes_per_inst": ps_["lanes_per_inst"],
                                     "stage_ms_per_step": ks[big]["ms"] / stat_steps, "achieved": ach / 1e9, "peak": issue_peak / 1e9, "unit": "Gwarp-inst/s",
                                     "frac": ach / issue_peak, "useful_lane_frac": ach / issue_peak * ps_["lanes_per_inst"] / 32.0,
                                     "source": os.path.relpath(prof_path, ROOT)}
        cpu = None
        relmse = None
        if not args.no_cpu_baseline and world == 1:
            relmse = relmse_check(ctx, args)
            v, cores, kind, sample = cpu_reference_sample(cfg, 15.0, args.cpu_spp)
            cpu = {"value": v, "unit": "paths/s", "cores": cores, "kind": kind, "sample": sample}
        out = {
            "metric": "paths/sec (guidedvolpathvspg render pass)", "value": value, "unit": "paths/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(cfg, args, world),
            "e2e": {"value": e2e_value, "unit": "paths/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps},
            "gpu_launches": int(round(cnt["kernel_launches"] * args.steps / stat_steps)), "clocks": clk, "roofline": roofline, "cpu_baseline": cpu, "relmse": relmse,
            "counters": {k: cnt[k] for k in ("paths", "density_queries", "dda_steps", "cache_lookups", "bvh_nodes", "shadow_rays", "wave_iterations")},
            "mean_radiance": float(film_host[..., :3].sum() / max(1.0, float(film_host[..., 3].sum()))),
        }
        print(json.dumps(out))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
