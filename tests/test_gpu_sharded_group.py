"""GPU, ONE device: the sharded key-frame map (BASELINE.json config 4) with its exchanges exercised for real — several shards
of one process on the same GPU (lo_sharded_create_group), each with its own stream. Exchange 1 (nearest stored point per
query across shards) runs inside the search kernel as atomicMin into the query's home shard + one cross-shard barrier + one
pull; exchange 2 (the 32 normal-equation sums) through the mailboxes inside the LM step kernel. No NCCL, so this runs on the
one-GPU box the driver tests on; between GPUs the same kernels write through NVLink peer memory (tests/test_gpu_sharded_native.py,
bench.py --gpus N). The result must equal single-table CloudMatcher::align and the CPU oracle."""
import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from util import planar_scene, pose_error

pytestmark = pytest.mark.gpu
GUESS = np.array([0.1, 0.05, -0.02, 0.9999619, 0, 0, 0.0087265], np.float32)


def scene(port, scans):
    kf = planar_scene(port, scans, 0)
    ds = port.grid(0.3, 1)
    ds.add_cloud(planar_scene(port, scans, 1))
    return kf, ds.get_cloud_xyz()


@pytest.mark.parametrize("world", [2, 3])
def test_shards_of_one_process_on_one_gpu(ctx, port, stream12, world):
    kf, q = scene(port, stream12[0])
    g = lo.VoxelGrid(ctx, 0.2, 20)
    g.addCloud(kf)
    want, stats = lo.CloudMatcher().align(g, q, GUESS, with_stats=True)
    ctxs = [lo.Context(0) for _ in range(world)]          # a stream per shard
    grp = lo.ShardedGroup(ctxs, 0.2, 20)
    grp.addCloud(kf)
    sizes = [gr.size() for gr in grp.grids]
    assert sum(sizes) == g.size() and sum(gr.numPoints() for gr in grp.grids) == g.numPoints()
    assert min(sizes) > 0.5 * g.size() / world             # hash-range partition: roughly even
    for run in range(3):                                   # repeated aligns: the buffers and sequence numbers carry over
        poses, st = grp.align(q, GUESS)
        for k in range(1, world):
            assert np.array_equal(poses[k].view(np.uint32), poses[0].view(np.uint32))     # every shard holds the same result, bit for bit
        assert st.outer_iterations == stats.outer_iterations
        assert list(st.pairs_log[:st.outer_iterations]) == list(stats.pairs_log[:stats.outer_iterations])   # same winners as one table
        assert st.lm_iterations == stats.lm_iterations and st.termination == stats.termination
        dt, dr = pose_error(poses[0], want)
        assert dt < 1e-6 and dr < 1e-6, (run, dt, dr)
    o = port.grid(0.2, 20)
    o.add_cloud(kf)
    ref_pose, _ = o.align(q, GUESS)
    dt, dr = pose_error(poses[0], ref_pose)
    assert dt < 1e-6 and dr < 1e-6, (dt, dr)
    info = grp.info(0)
    assert (info.world, info.fused_sum, info.fused_min, info.min_allreduces, info.sum_allreduces) == (world, 1, 1, 0, 0)
    assert info.min_exchanges_fused >= 3 * 5
    # other query counts in a row (fewer and more blocks; home ranges move), and the empty set = the prior-only solve
    for m in (1, 37, 1000):
        poses, st = grp.align(q[:m], GUESS)
        one, st1 = lo.CloudMatcher().align(g, q[:m], GUESS, with_stats=True)
        assert list(st.pairs_log[:st.outer_iterations]) == list(st1.pairs_log[:st1.outer_iterations]), m
        dt, dr = pose_error(poses[0], one)
        assert dt < 1e-6 and dr < 1e-6, (m, dt, dr)
    poses0, _ = grp.align(q[:0], GUESS)
    assert np.array_equal(poses0[0], lo.CloudMatcher().align(g, q[:0], GUESS))
    # a group is driven as a group
    with pytest.raises(lo.LoError):
        out = np.empty(7, np.float32)
        lo._check(grp.lib.lo_sharded_align_dev(grp.h[0], None, 0, lo._p(lo._pose7(GUESS)), 0.3, lo._p(out), None))
    grp.close()
    for c in ctxs:
        c.close()
