"""Instruction-count histogram of one kernel by source file:line (from `nvdisasm -g -c`), to find what bloats a kernel.
usage: sass_hist.py all.sass <kernel-substring> [top]"""
import sys, re, collections
path, key = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
cur = None; inside = False; hist = collections.Counter(); total = 0
for line in open(path, errors="ignore"):
    if line.startswith("//--------------------- .text."):
        inside = key in line
        continue
    if not inside:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        hist[cur] += 1; total += 1
print("total instructions", total, "=", total * 16 / 1024, "KB")
byfile = collections.Counter()
for (f, l), c in hist.items():
    byfile[f] += c
print("by file:", dict(byfile.most_common()))
for (f, l), c in hist.most_common(top):
    print(f"{c:6d}  {f}:{l}")
