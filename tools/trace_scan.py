"""Phase timeline of the per-scan kernel (LO_SCAN_TRACE=1): runs the bench stream for a few scans and aggregates the
block-0 clock deltas per tag. Tags 101-109: pipeline phases before align (prep, deskew, cell fill, curvature, planar, scatter, select, emit+bucket scan,
bucket placement); 0-9: align_device (see lo_align.cuh);
120-124: gate, cleanup, locate, chain, write."""
import os, subprocess, sys, collections, re
env = dict(os.environ, LO_SCAN_TRACE="1")
scans = int(sys.argv[1]) if len(sys.argv) > 1 else 30
code = f"""
import numpy as np, lidar_odometry_demo_b200 as lo
from bench import make_stream
ctx = lo.Context(0)
od = lo.LidarOdometry(ctx)
for sc in make_stream({scans}, 0):
    od.processCloud(sc)
"""
r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if r.returncode:
    print(r.stderr[-3000:]); sys.exit(1)
per_scan, cur = [], collections.OrderedDict()
for line in r.stderr.splitlines():
    m = re.match(r"trace (\w+) (-?\d+)", line)
    if not m:
        continue
    if m.group(1) == "end":
        cur["total"] = int(m.group(2)); per_scan.append(cur); cur = collections.OrderedDict()
    else:
        k = int(m.group(1)); cur.setdefault(k, [0, 0]); cur[k][0] += int(m.group(2)); cur[k][1] += 1
last = per_scan[len(per_scan) // 2:]
agg = collections.OrderedDict()
for sc in last:
    for k, v in sc.items():
        if k == "total":
            agg.setdefault(k, [0, 0]); agg[k][0] += v; agg[k][1] += 1
        else:
            agg.setdefault(k, [0, 0]); agg[k][0] += v[0]; agg[k][1] += v[1]
n = len(last)
print(f"{n} scans averaged; clocks per scan (count per scan)")
for k, v in agg.items():
    print(f"  tag {k}: {v[0] / n:10.0f} clk  ({v[1] / n:.1f}x, {v[0] / max(v[1], 1):.0f} each)")
