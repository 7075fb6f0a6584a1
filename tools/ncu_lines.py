#!/usr/bin/env python
"""Warp-stall samples of one kernel per CUDA-C source line: joins the SASS view of an ncu report
(`ncu -i X.ncu-rep --page source --csv`) with the line table of the cubin (`nvdisasm -g`). This ncu's command line
prints per-line metrics only for SASS, hence the join.
   cuobjdump -xelf all liblo_b200.so; nvdisasm -g lo_pipeline.sm_100a.cubin > all.sass
   python tools/ncu_lines.py scan1_source.csv all.sass _ZN2lo11scan_kernelENS_8ScanArgsE out.txt [top]"""
import collections
import csv
import os
import re
import sys

src, sass, kernel, out = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 45
line_of, cur, inside = {}, None, False
for l in open(sass, errors="replace"):
    if l.startswith(".text."):
        inside = l.strip() == f".text.{kernel}:"
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", l)
    if m and cur:
        line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src, newline="", errors="replace")))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
col = {c: i for i, c in enumerate(hdr)}
stalls = [c for c in hdr if c.startswith("stall_") and "Not Issued" not in c]
data = [r for r in rows[h + 1:] if len(r) >= len(hdr) and r[0].startswith("0x")]
base = min(int(r[0], 16) for r in data)
agg = collections.defaultdict(lambda: [0.0, 0.0, collections.Counter()])
for r in data:
    key = line_of.get(int(r[0], 16) - base, ("?", 0))
    a = agg[key]
    a[0] += float(r[col["Warp Stall Sampling (All Samples)"]] or 0)
    a[1] += float(r[col["Instructions Executed"]] or 0)
    for s in stalls:
        a[2][s] += float(r[col[s]] or 0)
total = sum(a[0] for a in agg.values()) or 1.0
text = {}
with open(out, "w") as f:
    f.write(f"# {kernel}: warp-stall samples per source line (total {total:.0f}), top {top}; columns: share, samples, warp instructions, file:line, two largest stall reasons\n")
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        why = ", ".join(f"{k[6:]} {100 * v / max(a[0], 1):.0f}%" for k, v in a[2].most_common(2))
        f.write(f"{100 * a[0] / total:5.1f}% {a[0]:8.0f} {a[1]:11.0f}  {key[0]}:{key[1]:<5d} {why}\n")
