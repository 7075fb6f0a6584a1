#!/usr/bin/env python
"""Where does the GPU pipeline leave the oracle on a stream? Prints every scan whose pose or counts differ.
   python tools/diag_parity.py <n_scans>"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

n = int(sys.argv[1])
scans, _, _ = S.vlp16_stream(n, workers=8)
import lidar_odometry_demo_b200 as lo  # noqa: E402
kind = "reference" if oracle.available("reference") else "port"
orc = oracle.load(kind)
ctx = lo.Context(0)
od, oc = lo.LidarOdometry(ctx), orc.odometry()
shown = 0
orders = [0, 0, 0]
for k, s in enumerate(scans):
    od.processCloud(s)
    c, extra = oc.process(s)
    a, b = od.getCurrentPose().astype(np.float64), oc.pose().astype(np.float64)
    dt = float(np.linalg.norm(a[:3] - b[:3]))
    st = od.stats()
    orders[st.update_order] += 1
    bad = dt > 1e-7 or (st.n_planar, st.n_update, st.n_query) != (c[1], c[2], c[3])
    if bad and shown < 12:
        shown += 1
        print(f"scan {k}: |dt| {dt:.3e}; GPU P/U/M {st.n_planar}/{st.n_update}/{st.n_query} outer {st.outer_iterations} unstable {st.unstable} kf {st.keyframe_voxels}; "
              f"oracle {c[1]}/{c[2]}/{c[3]} extra {extra}; gpu {a[:3]} oracle {b[:3]}", flush=True)
print("done; oracle", kind, "; update order: counting sort / replayed / approximate =", orders)
