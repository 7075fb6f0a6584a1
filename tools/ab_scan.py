#!/usr/bin/env python
"""A/B of two builds of liblo_b200.so on the same box: K scans through lo_odometry_submit_dev / lo_odometry_wait (queue kept
full), per-step CUDA events, with and without an L2 flush before every step. Plain ctypes (only the handful of entry points
every build has), so an older library can be timed next to the current one.
   python tools/ab_scan.py <libA.so> <libB.so> [scans]"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

libs, K, W = sys.argv[1:3], int(sys.argv[3]) if len(sys.argv) > 3 else 40, 10
scans, _, _ = S.vlp16_stream(W + K, workers=8)
import torch  # noqa: E402
dev = torch.device("cuda", 0)
d_scans = [torch.from_numpy(s.view(np.uint8).reshape(-1)).to(dev) for s in scans]
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def run(path, flush):
    L = C.CDLL(path)
    vp = C.c_void_p
    L.lo_ctx_create.argtypes, L.lo_ctx_stream.restype, L.lo_ctx_stream.argtypes = [C.c_int, C.POINTER(vp)], vp, [vp]
    L.lo_odometry_create.argtypes = [vp, vp, C.POINTER(vp)]
    L.lo_odometry_submit_dev.argtypes, L.lo_odometry_wait.argtypes = [vp, vp, C.c_size_t], [vp]
    L.lo_odometry_pose.argtypes, L.lo_params_default.argtypes = [vp, vp], [vp]
    L.lo_odometry_destroy.argtypes, L.lo_ctx_destroy.argtypes = [vp], [vp]
    ctx, od = vp(), vp()
    assert L.lo_ctx_create(0, C.byref(ctx)) == 0
    params = (C.c_uint8 * 64)()
    L.lo_params_default(params)
    assert L.lo_odometry_create(ctx, params, C.byref(od)) == 0
    stream = torch.cuda.ExternalStream(L.lo_ctx_stream(ctx), device=dev)
    for k in range(W):
        assert L.lo_odometry_submit_dev(od, d_scans[k].data_ptr(), len(scans[k])) == 0
        assert L.lo_odometry_wait(od) == 0
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    torch.cuda.synchronize()
    issued = 0
    for i in range(K):
        while issued < K and issued - i < 8:
            if flush:
                with torch.cuda.stream(stream):
                    flush_buf.fill_(1)
            ev[issued][0].record(stream)
            assert L.lo_odometry_submit_dev(od, d_scans[W + issued].data_ptr(), len(scans[W + issued])) == 0
            ev[issued][1].record(stream)
            issued += 1
        assert L.lo_odometry_wait(od) == 0
    torch.cuda.synchronize()
    ms = np.array([a.elapsed_time(b) for a, b in ev])
    pose = (C.c_float * 7)()
    L.lo_odometry_pose(od, pose)
    L.lo_odometry_destroy(od)
    L.lo_ctx_destroy(ctx)
    return ms, np.array(pose[:])


for rep in range(2):
    for flush in (True, False):
        for path in libs:
            ms, pose = run(path, flush)
            print(f"rep {rep} {'cold L2' if flush else 'warm L2'} {os.path.basename(path):22s} p50 {np.median(ms):.4f} ms  mean {ms.mean():.4f}  min {ms.min():.4f}  max {ms.max():.4f}  pose x {pose[0]:.6f}")
