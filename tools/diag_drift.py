"""Diagnostic (GPU box): where does GPU-vs-oracle pose drift over a long stream come from?
   a) full pipeline  b) planar entry (oracle deskew+classifier feed both)  c) per-scan stage mismatch counts."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import lidar_odometry_demo_b200 as lo
import oracle
from lidar_odometry_demo_b200 import synthetic as S

n = int(sys.argv[1]) if len(sys.argv) > 1 else 45
port = oracle.load("port")
ctx = lo.Context(0)
scans, pos, q = S.vlp16_stream(n)

def err(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    qa, qb = a[3:] / np.linalg.norm(a[3:]), b[3:] / np.linalg.norm(b[3:])
    return np.linalg.norm(a[:3] - b[:3]), 2 * np.arccos(min(1.0, abs(qa @ qb)))

od, oc = lo.LidarOdometry(ctx), port.odometry()
od2, oc2 = lo.LidarOdometry(ctx), port.odometry()
ident = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
for k, s in enumerate(scans):
    od.processCloud(s); c, _ = oc.process(s)
    e_full = err(od.getCurrentPose(), oc.pose())
    # planar entry: feed the oracle's planar cloud (from the oracle's own deskewed cloud) to both
    desk = oc.deskewed()
    planar, _ = port.classify(desk)
    od2.processPlanar(planar); c2, _ = oc2.process_planar(planar)
    e_pl = err(od2.getCurrentPose(), oc2.pose())
    st, st2 = od.stats(), od2.stats()
    # stage mismatch for this scan: GPU deskew+classify on the same input/pose as the oracle
    g_desk = od.getTempCloud()
    dmis = sum(int((g_desk[f].view(np.uint32) != desk[f].view(np.uint32)).sum()) for f in "xyz")
    print(f"{k:3d} full dt={e_full[0]:.2e} dr={e_full[1]:.2e} | planar-entry dt={e_pl[0]:.2e} dr={e_pl[1]:.2e} | "
          f"P {st.n_planar}/{c[1]} U {st.n_update}/{c[2]} M {st.n_query}/{c[3]} | pl U {st2.n_update}/{c2[2]} M {st2.n_query}/{c2[3]} "
          f"kf {st2.keyframe_voxels}/{len(oc2.keyframe(False))} | deskew coord mismatches {dmis}", flush=True)
