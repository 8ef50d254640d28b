"""Summarises an `ncu --metrics ... --csv` launch list (long format, one row per metric per launch) into per-kernel and
per-stage totals: launches, time, warp / thread instructions, DRAM bytes.  Written as JSON for profiles/ and read back
by bench.py (roofline.traffic, roofline.issue).
    python tools/ncu_metrics_summary.py gpurun_out/x.csv profiles/r02_ncu_metrics_C2.json "C2 1 render, spp 64" [spp of the profiled render] """
import csv, json, re, sys

STAGE = [("k_generate", "generate"), ("k_intersect", "intersect"), ("k_medium", "medium"), ("k_lookup<0", "volscatter"), ("k_volscatter", "volscatter"),
         ("k_surface", "surface"), ("k_lookup<1", "shade"), ("k_shade", "shade"), ("k_shadow", "shadow"), ("k_finish", "finish")]
METRICS = {"gpu__time_duration.sum": "ns", "smsp__inst_executed.sum": "warp_inst", "smsp__thread_inst_executed.sum": "thread_inst",
           "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write"}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6, "inst": 1.0, "": 1.0, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6}


def main(path, out, note, spp=0):
    rows = list(csv.reader(open(path, newline="")))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    ix = {k: i for i, k in enumerate(rows[h])}
    per = {}
    for r in rows[h + 1:]:
        if len(r) <= ix["Metric Value"] or r[ix["Metric Name"]] not in METRICS or not re.match(r"^(void )?(vspg::)?k_", r[ix["Kernel Name"]]):
            continue   # (workload builders may run torch kernels of their own: not part of the render)
        name = re.sub(r"^void ", "", r[ix["Kernel Name"]]).replace("vspg::", "")
        name = name.split("(vspg")[0].split("(DScene")[0].strip()
        name = re.sub(r"\(.*$", "", name) if not name.startswith("k_lookup") and "<" not in name else name.split("(const")[0]
        k = per.setdefault(name, {"launches": set(), "ns": 0.0, "warp_inst": 0.0, "thread_inst": 0.0, "dram_read": 0.0, "dram_write": 0.0})
        k["launches"].add(r[ix["ID"]])
        v = float(r[ix["Metric Value"]].replace(",", "")) * UNIT.get(r[ix["Metric Unit"]], 1.0)
        k[METRICS[r[ix["Metric Name"]]]] += v
    kernels, stages = {}, {}
    for name, k in sorted(per.items(), key=lambda kv: -kv[1]["ns"]):
        n = len(k["launches"])
        e = {"launches": n, "ms": k["ns"] / 1e6, "warp_inst": k["warp_inst"], "lanes_per_inst": k["thread_inst"] / max(1.0, k["warp_inst"]),
             "dram_read_bytes": k["dram_read"], "dram_write_bytes": k["dram_write"], "dram_bytes_per_launch": (k["dram_read"] + k["dram_write"]) / max(1, n)}
        kernels[name] = e
        st = next((s for pat, s in STAGE if name.startswith(pat)), "other")
        a = stages.setdefault(st, {"ms": 0.0, "warp_inst": 0.0, "thread_inst": 0.0, "dram_bytes": 0.0, "launches": 0})
        a["ms"] += e["ms"]; a["warp_inst"] += k["warp_inst"]; a["thread_inst"] += k["thread_inst"]; a["dram_bytes"] += k["dram_read"] + k["dram_write"]; a["launches"] += n
    for a in stages.values():
        a["lanes_per_inst"] = a["thread_inst"] / max(1.0, a["warp_inst"])
    tot = sum(a["ms"] for a in stages.values())
    doc = {"source": path, "note": note, "spp": spp, "what": "ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none (per-launch, serialised, cold cache: shares and counts are meaningful, absolute times are not bench values)",
           "total_ms_under_ncu": tot, "stages": stages, "kernels": kernels}
    json.dump(doc, open(out, "w"), indent=1)
    print(f"{'stage':12s} {'ms':>9s} {'share':>6s} {'Gwarp-inst':>11s} {'lanes':>6s} {'DRAM GB':>8s}")
    for s, a in sorted(stages.items(), key=lambda kv: -kv[1]["ms"]):
        print(f"{s:12s} {a['ms']:9.2f} {a['ms'] / tot:6.3f} {a['warp_inst'] / 1e9:11.3f} {a['lanes_per_inst']:6.1f} {a['dram_bytes'] / 1e9:8.2f}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "", int(sys.argv[4]) if len(sys.argv) > 4 else 0)
