#!/usr/bin/env python
"""Quick timing of the stand-alone fused NN+residual kernel on a key-frame map much larger than L2 (config 4 shape).
   python tools/bench_nn.py [map_points] [queries]"""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lidar_odometry_demo_b200 as lo  # noqa: E402
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith('--')]
n_map = int(args[0]) if len(args) > 0 else 12_000_000
n_q = int(args[1]) if len(args) > 1 else 262144
ctx = lo.Context(0)
dev = torch.device("cuda", 0)
stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
big = S.planar_patch_map(n_map, seed=777, extent=(400.0, 400.0, 40.0))
gb = lo.VoxelGrid(ctx, 0.2, 20, capacity_hint=n_map // 8)
for i in range(0, len(big), 2_000_000):
    gb.addCloud(big[i:i + 2_000_000])
rng = np.random.default_rng(5)
sel = rng.choice(len(big), n_q, replace=False)
qq = np.zeros(len(sel), S.XYZ)
qq["x"], qq["y"], qq["z"] = big["x"][sel] + 0.03, big["y"][sel] - 0.02, big["z"][sel] + 0.01
if "--sorted" in sys.argv:   # spatially coherent query order (what lo_nn_residual could do internally)
    k = (np.floor(qq["x"] / 1.6).astype(np.int64) << 40) | (np.floor(qq["y"] / 1.6).astype(np.int64) << 20) | np.floor(qq["z"] / 1.6).astype(np.int64)
    qq = qq[np.argsort(k, kind="stable")]
d_qq = torch.from_numpy(qq.view(np.uint8).reshape(-1)).to(dev)
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
ident = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
xi = np.array([1, 0, 0, 0, 0, 0, 0], np.float64)
zero3 = np.zeros(3)
neq = lo.NormalEq()
ts = []
for i in range(12):
    with torch.cuda.stream(stream):
        flush.fill_(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    rc = ctx.lib.lo_nn_residual_dev(gb.h, C.c_void_p(d_qq.data_ptr()), len(qq), ident.ctypes.data_as(C.c_void_p),
                                    xi.ctypes.data_as(C.c_void_p), zero3.ctypes.data_as(C.c_void_p), C.c_float(0.3), None)
    e1.record(stream)
    torch.cuda.synchronize()
    assert rc == 0
    if i >= 3:
        ts.append(e0.elapsed_time(e1) / 1e3)
ctx.lib.lo_nn_residual_dev(gb.h, C.c_void_p(d_qq.data_ptr()), len(qq), ident.ctypes.data_as(C.c_void_p), xi.ctypes.data_as(C.c_void_p),
                           zero3.ctypes.data_as(C.c_void_p), C.c_float(0.3), C.byref(neq))
t = float(np.mean(ts))
c = neq.candidates / len(qq)
alg = len(qq) * (464 + 16 * c)
peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
print(json.dumps({"ms": 1e3 * t, "min_ms": 1e3 * min(ts), "mpts_per_s": len(qq) / t / 1e6, "cand_per_query": c, "alg_GBps": alg / t / 1e9,
                  "frac_of_measured_hbm": alg / t / 1e9 / peak, "voxels": gb.size(), "points": gb.numPoints(), "valid": int(neq.n_valid),
                  "cost": neq.cost}))
