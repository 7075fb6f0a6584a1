#!/usr/bin/env python
"""Short driver for `ncu --set full`: a few scans through the pipeline (align_kernel launches) and the
stand-alone fused NN+residual kernel on a map much larger than L2. Not a benchmark."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lidar_odometry_demo_b200 as lo  # noqa: E402
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

n_map = int(sys.argv[1]) if len(sys.argv) > 1 else 12_000_000
ctx = lo.Context(0)
scans, _, _ = S.vlp16_stream(6)
od = lo.LidarOdometry(ctx)
for s in scans:
    od.processCloud(s)
print("pose", od.getCurrentPose(), "outer", od.stats().outer_iterations)

big = S.planar_patch_map(n_map, seed=777, extent=(400.0, 400.0, 40.0))
gb = lo.VoxelGrid(ctx, 0.2, 20, capacity_hint=n_map // 8)
for i in range(0, len(big), 2_000_000):
    gb.addCloud(big[i:i + 2_000_000])
rng = np.random.default_rng(5)
sel = rng.choice(len(big), 262144, replace=False)
qq = np.zeros(len(sel), S.XYZ)
qq["x"], qq["y"], qq["z"] = big["x"][sel] + 0.03, big["y"][sel] - 0.02, big["z"][sel] + 0.01
ident = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
x = np.array([1, 0, 0, 0, 0, 0, 0], np.float64)
for _ in range(4):
    cost, g, h, nv = gb.nnResidual(qq, ident, x, np.zeros(3))
print("nn_residual", cost, nv, gb.last_candidates / len(qq))
