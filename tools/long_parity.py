#!/usr/bin/env python
"""BASELINE.json config 2 at full length (1 000 scans of the synthetic VLP16 stream): the GPU pipeline against the CPU oracle
scan by scan — pose difference, planar / update / query counts, key-frame size. Prints a line every 50 scans and a JSON
summary.   python tools/long_parity.py [scans]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lidar_odometry_demo_b200 as lo  # noqa: E402
import oracle  # noqa: E402
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
kind = "reference" if oracle.available("reference") else "port"
orc = oracle.load(kind)
t0 = time.perf_counter()
scans, pos, q = S.vlp16_stream(n, workers=max(1, min(16, os.cpu_count() or 1)))
print(f"{n} scans generated in {time.perf_counter() - t0:.0f} s; oracle = {kind}", flush=True)
ctx = lo.Context(0)   # after the stream was generated: the generator forks worker processes
od, oc = lo.LidarOdometry(ctx), orc.odometry()
worst_t = worst_r = 0.0
count_mismatch = unstable = 0
orders = [0, 0, 0]
t_gpu = t_cpu = 0.0
for k, s in enumerate(scans):
    t1 = time.perf_counter()
    od.processCloud(s)
    t2 = time.perf_counter()
    c, _ = oc.process(s)
    t3 = time.perf_counter()
    t_gpu += t2 - t1
    t_cpu += t3 - t2
    a, b = od.getCurrentPose().astype(np.float64), oc.pose().astype(np.float64)
    dt = float(np.linalg.norm(a[:3] - b[:3]))
    qa, qb = a[3:] / np.linalg.norm(a[3:]), b[3:] / np.linalg.norm(b[3:])
    dr = float(2 * np.arccos(min(1.0, abs(qa @ qb))))
    worst_t, worst_r = max(worst_t, dt), max(worst_r, dr)
    st = od.stats()
    if (st.n_planar, st.n_update, st.n_query) != (c[1], c[2], c[3]):
        count_mismatch += 1
    unstable += int(st.unstable)
    orders[st.update_order] += 1
    if k % 50 == 49 or k == n - 1:
        print(f"scan {k:4d}: |dt| {dt:.2e} m |dr| {dr:.2e} rad (worst so far {worst_t:.2e} / {worst_r:.2e}); P/U/M {st.n_planar}/{st.n_update}/"
              f"{st.n_query} vs {c[1]}/{c[2]}/{c[3]}; key-frame voxels {st.keyframe_voxels}; x = {a[0]:.3f} m", flush=True)
print(json.dumps({"workload": f"config 2, {n} scans", "oracle": kind, "max_pose_err_m": worst_t, "max_pose_err_rad": worst_r,
                  "scans_with_count_mismatch": count_mismatch,
                  "update_order": {"counting_sort_exact": orders[0], "replayed": orders[1], "approximate": orders[2]}, "unstable_scans": unstable, "final_pose_gpu": [float(v) for v in a],
                  "final_pose_oracle": [float(v) for v in b], "gpu_wall_s": t_gpu, "cpu_wall_s": t_cpu, "speedup_wall_clock_blocking_pageable": t_cpu / t_gpu,
                  "keyframe": od.keyframeInfo().as_dict(),
                  "tolerance": "1e-4 m / 1e-4 rad per scan (north_star)"}))
