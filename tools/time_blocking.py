#!/usr/bin/env python
"""Wall clock of the blocking, processCloud-shaped call on pageable host scans, with the library's own host timers
(LO_B200_TIMING=1: staging upload / submit / wait for the pose, per call).   LO_B200_TIMING=1 python tools/time_blocking.py [scans]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
scans, _, _ = S.vlp16_stream(n, workers=8)
import lidar_odometry_demo_b200 as lo  # noqa: E402
ctx = lo.Context(0)
od = lo.LidarOdometry(ctx)
ts = []
for s in scans:
    t0 = time.perf_counter()
    od.processCloudHost(s.ctypes.data, len(s))
    ts.append(time.perf_counter() - t0)
ts = np.array(ts[50:]) * 1e3
print(f"blocking call, pageable host scans: mean {ts.mean():.4f} ms  p50 {np.median(ts):.4f}  p90 {np.percentile(ts, 90):.4f}  -> {1e3 / ts.mean():.0f} scans/s")
