#!/usr/bin/env python
"""Block-0 phase trace (LO_SCAN_TRACE=1) and CUDA-event time of the same blocking scans, for one build of the library.
   LO_SCAN_TRACE=1 python tools/ab_trace.py <lib.so> [scans]"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

path, K = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30
scans, _, _ = S.vlp16_stream(K, workers=8)
import torch  # noqa: E402
dev = torch.device("cuda", 0)
d_scans = [torch.from_numpy(s.view(np.uint8).reshape(-1)).to(dev) for s in scans]
L = C.CDLL(path)
vp = C.c_void_p
L.lo_ctx_create.argtypes, L.lo_ctx_stream.restype, L.lo_ctx_stream.argtypes = [C.c_int, C.POINTER(vp)], vp, [vp]
L.lo_odometry_create.argtypes = [vp, vp, C.POINTER(vp)]
L.lo_odometry_process_dev.argtypes, L.lo_params_default.argtypes = [vp, vp, C.c_size_t], [vp]
ctx, od = vp(), vp()
assert L.lo_ctx_create(0, C.byref(ctx)) == 0
params = (C.c_uint8 * 64)()
L.lo_params_default(params)
assert L.lo_odometry_create(ctx, params, C.byref(od)) == 0
stream = torch.cuda.ExternalStream(L.lo_ctx_stream(ctx), device=dev)
ms = []
for k in range(K):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    assert L.lo_odometry_process_dev(od, d_scans[k].data_ptr(), len(scans[k])) == 0
    e1.record(stream)
    torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
print("events ms (last half):", np.round(ms[K // 2:], 4).tolist(), "median", float(np.median(ms[K // 2:])))
