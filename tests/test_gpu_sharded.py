"""GPU: the sharded-map phases of the C ABI (lo_grid_set_shard, lo_shard_search_dev, lo_shard_evaluate_dev, lo_lm_*)
driven by lidar_odometry_demo_b200/sharded.py. Several ranks are emulated as several shards of one process on one
GPU (their kernels run back to back on one stream; no kernel waits on another)."""
import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from lidar_odometry_demo_b200 import sharded
from util import planar_scene, pose_error

pytestmark = pytest.mark.gpu
GUESS = np.array([0.1, 0.05, -0.02, 0.9999619, 0, 0, 0.0087265], np.float32)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_sharded_align_matches_single_gpu_align(ctx, port, stream12, world):
    scans = stream12[0]
    kf = planar_scene(port, scans, 0)
    ds = port.grid(0.3, 1)
    ds.add_cloud(planar_scene(port, scans, 1))
    q = ds.get_cloud_xyz()
    g = lo.VoxelGrid(ctx, 0.2, 20)
    g.addCloud(kf)
    want, stats = lo.CloudMatcher().align(g, q, GUESS, with_stats=True)
    shards = [sharded.GpuShard(lo, ctx, 0.2, 20, r, world) for r in range(world)]
    for s in shards:
        s.add_cloud(kf)
        s.set_queries(q)
    assert sum(s.grid.size() for s in shards) == g.size()          # the hash ranges partition the voxels
    assert sum(s.grid.numPoints() for s in shards) == g.numPoints()
    if world > 1:
        assert min(s.grid.size() for s in shards) > 0.5 * g.size() / world
    m = sharded.ShardedCloudMatcher(shards, sharded.TorchCollective(None), sharded.HostLM(ctx.lib), stream=shards[0].stream)
    pose, log = m.align(GUESS)
    assert len(log) == stats.outer_iterations
    assert [l[0] for l in log] == list(stats.pairs_log[:len(log)])   # identical correspondence counts per outer iteration
    dt, dr = pose_error(pose, want)
    assert dt < 1e-6 and dr < 1e-6, (dt, dr)
    # and against the CPU oracle
    o = port.grid(0.2, 20)
    o.add_cloud(kf)
    ref_pose, _ = o.align(q, GUESS)
    dt, dr = pose_error(pose, ref_pose)
    assert dt < 1e-6 and dr < 1e-6, (dt, dr)
