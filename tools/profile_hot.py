#!/usr/bin/env python
"""Short driver for `ncu --set full`: a few scans through the pipeline (scan_kernel), one stand-alone align
(align_kernel), the fused NN+residual kernels (tile search and exhaustive) on a map much larger than L2, three 128-beam
scans of config 3 (scan_kernel_large) and one device-resident sharded align with one shard on the big map
(shard_search_ctl / shard_evaluate_ctl / shard_step kernels). Not a benchmark."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lidar_odometry_demo_b200 as lo  # noqa: E402
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

n_map = int(sys.argv[1]) if len(sys.argv) > 1 else 12_000_000
ctx = lo.Context(0)
scans, _, _ = S.vlp16_stream(8)
od = lo.LidarOdometry(ctx)
for s in scans:
    od.processCloud(s)
print("pose", od.getCurrentPose(), "outer", od.stats().outer_iterations)

# CloudMatcher::align alone, on the pipeline's key frame and the last scan's matching set
kf = od.getKeyFrameCloudWithNormals()
g = lo.VoxelGrid(ctx, od.params.keyframe_voxel_size, int(od.params.keyframe_max_points_cnt))
g.addCloud(kf)
ident = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
planar = ctx.range_filter(ctx.classify(ctx.deskew(scans[-1], ident, ident)), od.params.lidar_min_range, od.params.lidar_max_range)
ds = lo.VoxelGrid(ctx, od.params.keyframe_matching_voxel_size, 1)
ds.addCloud(planar)
q = ds.getCloudWithoutNormals()
for _ in range(2):
    pose, st = lo.CloudMatcher().align(g, q, od.getCurrentPose(), with_stats=True)
print("align", pose, st.outer_iterations)

big = S.planar_patch_map(n_map, seed=777, extent=(400.0, 400.0, 40.0))
gb = lo.VoxelGrid(ctx, 0.2, 20, capacity_hint=n_map // 8)
for i in range(0, len(big), 2_000_000):
    gb.addCloud(big[i:i + 2_000_000])
rng = np.random.default_rng(5)
sel = rng.choice(len(big), 262144, replace=False)
qq = np.zeros(len(sel), S.XYZ)
qq["x"], qq["y"], qq["z"] = big["x"][sel] + 0.03, big["y"][sel] - 0.02, big["z"][sel] + 0.01
x = np.array([1, 0, 0, 0, 0, 0, 0], np.float64)
for mode in ("tile", "exhaustive"):
    if mode == "exhaustive":
        os.environ["LO_B200_NN_EXHAUSTIVE"] = "1"
    for _ in range(3):
        cost, grad, hess, nv = gb.nnResidual(qq, ident, x, np.zeros(3))
    print("nn_residual", mode, cost, nv, gb.last_candidates / len(qq))

# config 3: 128 beams x 2048 columns through the per-scan kernel's large instantiation
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_c3 import c3_params, c3_stream  # noqa: E402
od3 = lo.LidarOdometry(ctx, c3_params(lo))
for s in c3_stream(3):
    od3.processCloud(s)
print("config 3", len(s), od3.getCurrentPose(), od3.stats().outer_iterations)

# the device-resident sharded matcher, one shard (no NCCL): same map and queries as the NN kernels above
m = lo.ShardedCloudMatcher(ctx, None, 0, 1, 0.2, 20, capacity_hint=n_map // 8)
for i in range(0, len(big), 2_000_000):
    m.grid.addCloud(big[i:i + 2_000_000])
pose, st = m.align(qq, ident, with_stats=True)
print("sharded align (1 shard)", pose, st.outer_iterations, m.info().evaluations)
