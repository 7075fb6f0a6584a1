#!/usr/bin/env python
"""Scan 251 of the 440-scan stream: are the two key frames equal before it, and what do the two aligns do on equal inputs?"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import oracle  # noqa: E402
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402
from util import sort_rows  # noqa: E402

n, stop = 440, int(sys.argv[1]) if len(sys.argv) > 1 else 251
scans, _, _ = S.vlp16_stream(n, workers=8)
import lidar_odometry_demo_b200 as lo  # noqa: E402
orc = oracle.load("reference" if oracle.available("reference") else "port")
ctx = lo.Context(0)
od, oc = lo.LidarOdometry(ctx), orc.odometry()
for k in range(stop):
    od.processCloud(scans[k])
    oc.process(scans[k])
    if k >= stop - 3:
        a, b = od.getKeyFrameCloudWithNormals(), oc.keyframe_pn()
        same = len(a) == len(b) and np.array_equal(sort_rows(a), sort_rows(b))
        print(f"after scan {k}: key-frame points GPU {len(a)} oracle {len(b)} equal as multisets: {same}; pose equal: "
              f"{np.array_equal(od.getCurrentPose().view(np.uint32), oc.pose().view(np.uint32))}", flush=True)
        if not same and len(a) == len(b):
            sa, sb = sort_rows(a), sort_rows(b)
            diff = np.flatnonzero((sa != sb).any(1))
            print("  differing rows:", len(diff), "first", diff[:5])
# the scan itself, stage by stage through the ABI on the oracle's key frame (exact robin order) and the oracle's align
kf = oc.keyframe_pn()
prev_pose = oc.pose()
g = lo.VoxelGrid(ctx, 0.2, 20, capacity_hint=1 << 18)
g.addCloud(kf)                      # NB: slot order inside a voxel = iteration order of the oracle's map here, not arrival order
o = orc.grid(0.2, 20)
o.add_cloud(kf)
sc = scans[stop]
od.processCloud(sc)
c, extra = oc.process(sc)
print("scan", stop, "GPU pose", od.getCurrentPose(), "oracle", oc.pose(), "counts", c)
