#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (profiles/*_ncu_launches_summary.txt).
   python tools/ncu_launch_summary.py gpurun_out/launches.csv profiles/r1_ncu_launches_summary.txt"""
import csv
import re
import sys
from collections import defaultdict

src, out = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
col = {name: k for k, name in enumerate(rows[0])}
total, count = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    if r[col["Metric Name"]] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r[col["Kernel Name"]])[:110]
    scale = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r[col["Metric Unit"]], 1.0)
    total[name] += float(r[col["Metric Value"]].replace(",", "")) * scale
    count[name] += 1
grand = sum(total.values()) or 1.0
with open(out, "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none  python bench.py --steps 5 --warmup 3\n"
            "# per-launch times under ncu are cold-cache and serialised; what must agree with bench.py is the SHARE of the step.\n"
            "# One timed step = one lo::scan_kernel launch + the lo::scan_replay_kernel queued behind it, which returns at once unless\n"
            "# the scan needs the exact robin_map replay (about 1 scan in 25; those launches take ~0.4 ms and make up its share here).\n"
            "# gpu_launches = 2 x steps (+ table rebuilds: none). The other kernels belong to bench.py's untimed sections: L2 flush\n"
            "# fills (at::FillFunctor), the roofline section (align_kernel, nn_residual*_kernel, grid_* inserts of the 12 M-point map),\n"
            "# odometry construction.\n")
    f.write(f"{'launches':>8s} {'total_ns':>14s} {'share':>7s}  kernel\n")
    for name in sorted(total, key=total.get, reverse=True):
        f.write(f"{count[name]:8d} {total[name]:14.1f} {100 * total[name] / grand:6.1f}%  {name}\n")
    k = "lo::scan_kernel"
    if k in total:
        f.write(f"# scan_kernel: {count[k]} launches, {total[k] / count[k] / 1e3:.1f} us per launch under ncu; bench.py (CUDA events, L2 flushed): "
                "ms_per_step of the bench line of the same round\n")
