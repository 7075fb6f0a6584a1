#!/usr/bin/env python
"""BASELINE.json config 3: 128-beam x 2048 = 262 144 returns per scan, scene extent x2, keyframe voxel 0.1, update /
matching voxel 0.05. Parity against the CPU oracle on the first scans and device time per scan.
   python tools/bench_c3.py [scans] [--no-oracle]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402


def c3_stream(n_scans):
    lidar = S.LidarModel.beam128()
    pos, q, _ = S.trajectory(n_scans + 1, 99)
    scene = S.Scene(4321, length=max(120.0, pos[-1, 0] + 280.0), scale=2.0)
    return [S.generate_scan(scene, lidar, pos[k], q[k], pos[k + 1], q[k + 1], 4321 + k, t_offset=0.1 * k) for k in range(n_scans)]


def c3_params(lo):
    p = lo.Params()
    lo.load_library().lo_params_default(__import__("ctypes").byref(p))
    p.keyframe_voxel_size, p.keyframe_update_voxel_size, p.keyframe_matching_voxel_size = 0.1, 0.05, 0.05
    p.keyframe_max_points_cnt = 20
    return p


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    n = int(args[0]) if args else 6
    t0 = time.perf_counter()
    scans = c3_stream(n)
    print(f"{n} scans of {len(scans[0])} points generated in {time.perf_counter() - t0:.1f} s", file=sys.stderr)
    import torch
    import lidar_odometry_demo_b200 as lo
    ctx = lo.Context(0)
    # sized once for this configuration (0.1 m key-frame voxels, 262 144-point scans): nothing grows inside the timed calls
    od = lo.LidarOdometry(ctx, c3_params(lo), keyframe_voxels=1 << 21, scan_points=1 << 18)
    oc = None
    if "--no-oracle" not in sys.argv:
        import oracle
        orc = oracle.load("reference" if oracle.available("reference") else "port")
        op = orc.default_params()
        op.keyframe_voxel_size, op.keyframe_update_voxel_size, op.keyframe_matching_voxel_size = 0.1, 0.05, 0.05
        op.keyframe_max_points_cnt = 20
        oc = orc.odometry(op)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", 0))
    d_scans = [torch.from_numpy(s.view(np.uint8).reshape(-1)).cuda() for s in scans]
    torch.cuda.synchronize()
    worst, ms, t_cpu = (0.0, 0.0), [], 0.0
    orders = [0, 0, 0]
    kf_equal = 0
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
    from util import sort_rows
    for k, sc in enumerate(scans):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        od.processCloudDevice(d_scans[k].data_ptr(), len(sc))
        e1.record(stream)
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
        st = od.stats()
        orders[st.update_order] += 1
        if oc is not None:
            t1 = time.perf_counter()
            counts, _ = oc.process(sc)
            t_cpu += time.perf_counter() - t1
            a, b = od.getCurrentPose().astype(np.float64), oc.pose().astype(np.float64)
            dt = float(np.linalg.norm(a[:3] - b[:3]))
            qa, qb = a[3:] / np.linalg.norm(a[3:]), b[3:] / np.linalg.norm(b[3:])
            dr = float(2 * np.arccos(min(1.0, abs(qa @ qb))))
            worst = (max(worst[0], dt), max(worst[1], dr))
            assert (st.n_planar, st.n_update, st.n_query) == (counts[1], counts[2], counts[3]), (k, st.n_planar, st.n_update, st.n_query, counts)
            ka, kb = od.getKeyFrameCloudWithNormals(), oc.keyframe_pn()   # the key frame as a multiset of (point, normal) bit patterns
            kf_equal += int(len(ka) == len(kb) and np.array_equal(sort_rows(ka), sort_rows(kb)))
        print(k, f"{ms[-1]:.3f} ms", st.n_planar, st.n_update, st.n_query, st.outer_iterations, st.keyframe_voxels, file=sys.stderr)
    print(json.dumps({"workload": "config 3: 128 x 2048 returns per scan, voxels 0.1 / 0.05 / 0.05", "scans": n, "points_per_scan": len(scans[0]),
                      "update_order": {"counting_sort_exact": orders[0], "replayed": orders[1], "approximate": orders[2]},
                      "ms_per_scan_after_first": float(np.mean(ms[1:])), "scans_per_s": 1e3 / float(np.mean(ms[1:])),
                      "cpu_oracle_scans_per_s": (n / t_cpu) if oc is not None else None,
                      "max_pose_err_m": worst[0], "max_pose_err_rad": worst[1],
                      "scans_with_keyframe_identical_to_the_oracle": kf_equal if oc is not None else None, "kernel_launches": int(ctx.kernel_launches),
                      "ms_per_scan": [float(v) for v in ms], "keyframe": od.keyframeInfo().as_dict()}))
