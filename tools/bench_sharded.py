#!/usr/bin/env python
"""BASELINE.json config 4: a large synthetic key-frame map (default 50 M points) sharded by voxel-hash range over the
ranks of a torchrun job, 262 144 replicated queries, CloudMatcher::align with an NCCL all-reduce(MIN) of the packed
correspondence keys and an all-reduce(SUM) of the 28 normal-equation sums per evaluation.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/bench_sharded.py [--points 50000000] [--queries 262144] [--repeat 5]
Prints one JSON line on rank 0 (device time = max over ranks): the Python-orchestrated path (per-rank phases + torch.distributed
collectives + host LM) and, under "native", the device-resident matcher of the library (lo_sharded_*: NCCL MIN exchange, the
32-double sum through peer memory inside the LM step kernel, and the same with ncclAllReduce for that sum)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lidar_odometry_demo_b200 as lo  # noqa: E402
from lidar_odometry_demo_b200 import sharded, synthetic as S  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--points", type=int, default=50_000_000)
ap.add_argument("--queries", type=int, default=262144)
ap.add_argument("--repeat", type=int, default=5)
args = ap.parse_args()

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = lo.Context(local)
shard = sharded.GpuShard(lo, ctx, 0.2, 20, rank, world, capacity_hint=args.points // (6 * world))
uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
if world > 1:
    if rank == 0:
        uid = torch.from_numpy(lo.sharded_unique_id()).cuda()
    dist.broadcast(uid, 0)
native = lo.ShardedCloudMatcher(ctx, uid.cpu().numpy() if world > 1 else None, rank, world, 0.2, 20, capacity_hint=args.points // (6 * world))

# the map: planar patches in a 1 km x 1 km x 40 m volume, generated in chunks, identical on every rank
t0 = time.perf_counter()
chunk, made, qsel = 2_000_000, 0, []
rng = np.random.default_rng(4242)
k = 0
while made < args.points:
    n = min(chunk, args.points - made)
    pn = S.planar_patch_map(n, seed=777 + k, extent=(1000.0, 1000.0, 40.0))
    shard.add_cloud(pn)
    native.grid.addCloud(pn)
    take = rng.choice(len(pn), max(1, args.queries * n // args.points), replace=False)
    qsel.append(np.stack([pn["x"][take], pn["y"][take], pn["z"][take]], 1))
    made += len(pn)
    k += 1
q3 = np.concatenate(qsel)[:args.queries]
# queries: map points moved by a small pose (the guess is the identity, so align has to recover it). SURVEY 8(d) asks for
# 0.1 m / 1 deg, but 1 deg about the centre of a 1 km map moves the rim by 8.7 m, far outside the 0.3 m matching radius;
# 0.01 deg keeps every displacement below 0.2 m.
ang = np.deg2rad(0.01)
Rz = np.array([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]])
c = np.array([500.0, 500.0, 20.0])
q3 = ((q3 - c) @ Rz.T + c + np.array([0.06, -0.05, 0.06])).astype(np.float32)
queries = S.to_xyz(q3)
shard.set_queries(queries)
build_s = time.perf_counter() - t0

coll = sharded.TorchCollective(dist if world > 1 else None)
guess = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
stream = shard.stream
times, pose, log, evals = [], None, None, 0
for it in range(args.repeat + 1):
    m = sharded.ShardedCloudMatcher([shard], coll, sharded.HostLM(ctx.lib), stream=stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    pose, log = m.align(guess)
    e1.record(stream)
    torch.cuda.synchronize()
    if it > 0:
        times.append(e0.elapsed_time(e1) / 1e3)
    evals = m.evaluations
# the same align through lo_sharded_align_dev (queries already resident in HBM: shard.q)
native_out = {}
modes = (("fused_peer_memory_sum", 1), ("nccl_sum", 0)) if world > 1 else (("one_shard", 0),)
for name, fused in modes:
    if world > 1:
        try:
            native.setFusedSum(fused)
        except lo.LoError as e:
            native_out[name] = {"unavailable": str(e)}
            continue
    nt = []
    for it in range(args.repeat + 1):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        s0 = native.info().host_syncs
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        e0.record(stream)
        npose, nstats = native.alignDevice(shard.q.data_ptr(), len(queries), guess)
        e1.record(stream)
        torch.cuda.synchronize()
        w1 = time.perf_counter()
        if it > 0:
            nt.append((e0.elapsed_time(e1) / 1e3, w1 - w0))
    tt = torch.tensor([float(np.mean([a for a, _ in nt])), float(np.mean([b for _, b in nt]))], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    info = native.info()
    native_out[name] = {"align_s": float(tt[0]), "align_wall_s": float(tt[1]), "outer_iterations": int(nstats.outer_iterations),
                        "evaluations": int(info.evaluations), "host_syncs_per_align": int(info.host_syncs - s0), "final_cost": float(nstats.final_cost),
                        "pose": [float(v) for v in npose], "pose_equals_python_path": bool(np.array_equal(npose, pose)),
                        "nccl_version": int(info.nccl_version), "fused_sum": int(info.fused_sum)}
# phase timings of one search + one evaluation (device, this rank)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
ev[0].record(stream)
loc = shard.search(guess)
ev[1].record(stream)
x0 = np.array([1, 0, 0, 0, 0, 0, 0], np.float64)
shard.evaluate(loc, x0)
ev[2].record(stream)
torch.cuda.synchronize()
t = torch.tensor([float(np.mean(times)), ev[0].elapsed_time(ev[1]) / 1e3, ev[1].elapsed_time(ev[2]) / 1e3], dtype=torch.float64, device="cuda")
sizes = torch.tensor([shard.grid.size(), shard.grid.numPoints()], dtype=torch.int64, device="cuda")
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(sizes)
if rank == 0:
    t_align, t_search, t_eval = [float(v) for v in t.tolist()]
    outers = len(log)
    print(json.dumps({
        "workload": f"config 4: {int(sizes[1])} key-frame points in {int(sizes[0])} voxels sharded by voxel-hash range over {world} GPU(s), "
                    f"{len(queries)} replicated queries, CloudMatcher::align",
        "n_gpus": world, "align_s": t_align, "outer_iterations": outers, "evaluations": evals,
        "allreduces_per_align": outers + evals, "queries_per_s_through_align": len(queries) * outers / t_align,
        "search_ms_per_pass_max_rank": 1e3 * t_search, "evaluate_ms_per_pass_max_rank": 1e3 * t_eval,
        "native": native_out, "valid_pairs_last": log[-1][0], "final_cost": log[-1][3], "pose": [float(v) for v in pose], "map_build_s": build_s,
        "expected_pose_note": "queries were moved by t=(0.06,-0.05,0.06) + 0.01 deg about z at (500,500,20); align returns the inverse motion"}))
if world > 1:
    dist.destroy_process_group()
