#!/usr/bin/env python
"""Wall time of lo_robin_iteration_order (one block: H2D of n hashes + the block replay + D2H of n indices) for a few n.
   python tools/time_robin.py"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lidar_odometry_demo_b200 as lo  # noqa: E402

ctx = lo.Context(0)
rng = np.random.default_rng(5)
for n in (16, 1024, 4096, 8192, 8193, 13000, 16384):
    h = rng.choice(1 << 22, n, replace=False).astype(np.uint32)
    ts = []
    for _ in range(30):
        t0 = time.perf_counter()
        ctx.robin_iteration_order(h)
        ts.append(time.perf_counter() - t0)
    print(f"n = {n:6d}: min {min(ts) * 1e6:8.1f} us   median {sorted(ts)[len(ts) // 2] * 1e6:8.1f} us", flush=True)
