#!/usr/bin/env python
"""Top source lines of one kernel by warp-stall samples, from `ncu -i X.ncu-rep --page source --csv` (CUDA-C view, needs
-lineinfo and --import-source on).   python tools/ncu_hot_lines.py source.csv out.txt [top]"""
import csv
import sys

src, out = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
rows = list(csv.reader(open(src, newline="", errors="replace")))
hdr_at = next(i for i, r in enumerate(rows) if any("Source" == c.strip() for c in r))
hdr = [c.strip() for c in rows[hdr_at]]
col = {c: i for i, c in enumerate(hdr)}
samp = next((c for c in hdr if c.startswith("# Samples") or c == "Warp Stall Sampling (All Samples)"), None) or next(c for c in hdr if "Samples" in c)
inst = next((c for c in hdr if c.startswith("Instructions Executed")), None)
data = []
for r in rows[hdr_at + 1:]:
    if len(r) < len(hdr):
        continue
    try:
        s = float(r[col[samp]].replace(",", "") or 0)
    except ValueError:
        continue
    n = r[col[inst]].replace(",", "") if inst else ""
    data.append((s, r[col["#"]] if "#" in col else "", n, r[col["Source"]].strip()))
total = sum(d[0] for d in data) or 1.0
data.sort(reverse=True)
with open(out, "w") as f:
    f.write(f"# {src}: top {top} source lines by '{samp}' (total {total:.0f}); columns: share, samples, line, instructions executed, source\n")
    for s, line, n, text in data[:top]:
        f.write(f"{100 * s / total:5.1f}% {s:9.0f} {line:>6s} {n:>12s}  {text[:150]}\n")
