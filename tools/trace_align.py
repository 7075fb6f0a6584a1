"""Developer aid (GPU box): LO_ALIGN_TRACE=1 python tools/trace_align.py -> per-phase SM-clock deltas of align_kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import lidar_odometry_demo_b200 as lo
import oracle
from lidar_odometry_demo_b200 import synthetic as S
port = oracle.load("port")
ctx = lo.Context(0)
scans, _, _ = S.vlp16_stream(40)
od = lo.LidarOdometry(ctx)
for s in scans:
    od.processCloud(s)
kf = od.getKeyFrameCloudWithNormals()
g = lo.VoxelGrid(ctx, 0.2, 20, capacity_hint=len(kf) // 4)
g.addCloud(kf)
ident = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
planar = ctx.range_filter(ctx.classify(ctx.deskew(scans[-1], ident, ident)), 4.0, 80.0)
ds = lo.VoxelGrid(ctx, 0.3, 1)
ds.addCloud(planar)
q = ds.getCloudWithoutNormals()
pose = od.getCurrentPose()
for i in range(3):
    print("---- run", i, "queries", len(q), file=sys.stderr)
    out, st = lo.CloudMatcher().align(g, q, pose, with_stats=True)
print("outer", st.outer_iterations, "lm", st.lm_iterations)
