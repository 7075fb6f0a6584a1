"""CPU suite: host-side logic of the sharded-map path (lidar_odometry_demo_b200/sharded.py) with world_size-2 gloo:
two processes, each holding the voxels of its hash range (oracle-backed stand-in shards), must reproduce the
unsharded oracle align."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import lidar_odometry_demo_b200 as lo
from lidar_odometry_demo_b200 import sharded
from lidar_odometry_demo_b200 import synthetic as S
from util import planar_scene, pose_error

GUESS = np.array([0.1, 0.05, -0.02, 0.9999619, 0, 0, 0.0087265], np.float32)


def _scene():
    import oracle
    port = oracle.load("port")
    scans, _, _ = S.vlp16_stream(2)
    kf = planar_scene(port, scans, 0)
    ds = port.grid(0.3, 1)
    ds.add_cloud(planar_scene(port, scans, 1))
    return port, kf, ds.get_cloud_xyz()


def _worker(rank, world, port_no, out):
    from sharded_util import OracleShard
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    port, kf, q = _scene()
    shard = OracleShard(port, 0.2, 20, rank, world)
    shard.add_cloud(kf)
    shard.set_queries(q)
    m = sharded.ShardedCloudMatcher([shard], sharded.TorchCollective(dist), sharded.HostLM(lo.load_library()))
    pose, log = m.align(GUESS)
    sizes = torch.tensor([shard.grid.size()], dtype=torch.int64)
    dist.all_reduce(sizes)
    if rank == 0:
        out.put((pose, [l[0] for l in log], int(sizes[0]), m.evaluations))
    dist.destroy_process_group()


def test_two_rank_gloo_matches_unsharded_oracle():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port_no = s.getsockname()[1]
    ctxm = mp.get_context("spawn")
    out = ctxm.Queue()
    procs = [ctxm.Process(target=_worker, args=(r, 2, port_no, out)) for r in range(2)]
    for p in procs:
        p.start()
    pose, pairs, voxels, evals = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    port, kf, q = _scene()
    g = port.grid(0.2, 20)
    g.add_cloud(kf)
    want, log = g.align(q, GUESS)
    assert voxels == g.size()                       # the two hash ranges partition the map
    assert pairs == [int(k) for k in log[:, 0]]     # same correspondences in every outer iteration
    dt, dr = pose_error(pose, want)
    assert dt < 1e-6 and dr < 1e-6, (dt, dr)
    assert len(log) <= evals <= 6 * len(log)        # 1 + <=5 evaluations (2 all-reduces each) per outer iteration


def test_single_process_two_local_shards():
    from sharded_util import OracleShard
    port, kf, q = _scene()
    shards = [OracleShard(port, 0.2, 20, r, 2) for r in range(2)]
    for s in shards:
        s.add_cloud(kf)
        s.set_queries(q)
    m = sharded.ShardedCloudMatcher(shards, sharded.TorchCollective(None), sharded.HostLM(lo.load_library()))
    pose, _ = m.align(GUESS)
    g = port.grid(0.2, 20)
    g.add_cloud(kf)
    want, _ = g.align(q, GUESS)
    dt, dr = pose_error(pose, want)
    assert dt < 1e-6 and dr < 1e-6
