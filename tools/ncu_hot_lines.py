"""Maps the per-instruction counters of an `ncu --page source --csv` export onto source lines, using nvdisasm's line info.
    cuobjdump -xelf all libvspg.so; nvdisasm --print-line-info vspg_ctx.sm_100a.cubin > /tmp/lib.disasm
    python tools/ncu_hot_lines.py gpurun_out/x_source.csv.gz /tmp/lib.disasm "k_volscatter<(int)8>" _ZN4vspg12k_volscatterILi8E [nth]"""
import collections, csv, gzip, re, sys


def main(src_csv, disasm, kname, section, nth=1, top=40):
    f = gzip.open(src_csv, "rt") if src_csv.endswith(".gz") else open(src_csv)
    seen, rows, cap, hdr = 0, [], False, None
    for row in csv.reader(f):
        if not row:
            continue
        if row[0] == "Kernel Name":
            if cap:
                break
            if kname in row[1]:
                seen += 1
                cap = seen == nth
            continue
        if not cap:
            continue
        if row[0] == "Address":
            hdr = {h: i for i, h in enumerate(row)}
            continue
        rows.append(row)
    lines = open(disasm).read().split("\n")
    start = next(i for i, l in enumerate(lines) if l.strip().startswith(".section") and (".text." + section) in l)
    cur, inst = None, []
    for l in lines[start + 1:]:
        if l.strip().startswith(".section"):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            inst.append((cur, m.group(2)))
    print(len(rows), "profiled instructions,", len(inst), "disassembled")
    agg, aggs, stall = collections.Counter(), collections.Counter(), collections.Counter()
    tot = tots = 0
    for i in range(min(len(inst), len(rows))):
        ex, smp = int(rows[i][hdr["Instructions Executed"]]), int(rows[i][hdr["# Samples"]])
        agg[inst[i][0]] += ex; aggs[inst[i][0]] += smp; tot += ex; tots += smp
        for h, j in hdr.items():
            if h.startswith("stall_") and not h.endswith("(Not Issued)"):
                try: stall[h] += int(rows[i][j])
                except ValueError: pass
    print("warp instructions", tot, "samples", tots)
    st = sum(stall.values())
    print("stalls:", " ".join("%s %.0f%%" % (h[6:], 100 * v / st) for h, v in stall.most_common(7)))
    byfile = collections.Counter()
    for k, v in agg.items():
        byfile[k[0] if k else None] += v
    print("by file:", " ".join("%s %.1f%%" % (k, 100 * v / tot) for k, v in byfile.most_common(8)))
    for k, v in agg.most_common(top):
        print(k, "inst %.1f%%" % (100 * v / tot), "samples %.1f%%" % (100 * aggs[k] / max(1, tots)))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4], int(sys.argv[5]) if len(sys.argv) > 5 else 1)
