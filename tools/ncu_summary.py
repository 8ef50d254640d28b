#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel markdown table.

    python tools/ncu_summary.py gpurun_out/launches.csv > profiles/rNN_launch_summary.md
"""
import csv
import re
import sys
from collections import defaultdict


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        rows.append((name, ns, r["Block Size"], r["Grid Size"]))
    agg = defaultdict(lambda: [0, 0.0, set()])
    for name, ns, b, g in rows:
        a = agg[name]
        a[0] += 1
        a[1] += ns
        a[2].add(b)
    tot = sum(a[1] for a in agg.values())
    print(f"launches: {len(rows)}, total {tot / 1e6:.2f} ms (ncu, serialised, cold cache)\n")
    print("| kernel | launches | sum ms | avg us | share | block |")
    print("|---|---|---|---|---|---|")
    for name, (n, ns, bs) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| {name} | {n} | {ns / 1e6:.2f} | {ns / n / 1e3:.1f} | {100 * ns / tot:.1f}% | {' '.join(sorted(bs))} |")


if __name__ == "__main__":
    main(sys.argv[1])
