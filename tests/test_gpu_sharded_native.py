"""GPU: the device-resident sharded matcher with the collectives inside the library (lo_sharded_*, csrc/lo_sharded.cu):
world = 1 on one GPU against single-table lo_align and the CPU oracle; world = 2 as two processes on two GPUs (NCCL MIN
exchange + the peer-memory sum exchange, and the ncclAllReduce variant) — skipped on a one-GPU box."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from util import planar_scene, pose_error

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GUESS = np.array([0.1, 0.05, -0.02, 0.9999619, 0, 0, 0.0087265], np.float32)


def scene(port, scans):
    kf = planar_scene(port, scans, 0)
    ds = port.grid(0.3, 1)
    ds.add_cloud(planar_scene(port, scans, 1))
    return kf, ds.get_cloud_xyz()


def test_world_one_matches_align_and_oracle(ctx, port, stream12):
    kf, q = scene(port, stream12[0])
    g = lo.VoxelGrid(ctx, 0.2, 20)
    g.addCloud(kf)
    want, stats = lo.CloudMatcher().align(g, q, GUESS, with_stats=True)
    m = lo.ShardedCloudMatcher(ctx, None, 0, 1, 0.2, 20)
    m.grid.addCloud(kf)
    assert (m.grid.size(), m.grid.numPoints()) == (g.size(), g.numPoints())
    pose, st = m.align(q, GUESS, with_stats=True)
    assert st.outer_iterations == stats.outer_iterations
    assert list(st.pairs_log[:st.outer_iterations]) == list(stats.pairs_log[:stats.outer_iterations])
    assert st.lm_iterations == stats.lm_iterations and st.termination == stats.termination
    dt, dr = pose_error(pose, want)
    assert dt < 1e-6 and dr < 1e-6, (dt, dr)
    o = port.grid(0.2, 20)
    o.add_cloud(kf)
    ref_pose, _ = o.align(q, GUESS)
    dt, dr = pose_error(pose, ref_pose)
    assert dt < 1e-6 and dr < 1e-6, (dt, dr)
    info = m.info()
    assert (info.world, info.fused_sum, info.fused_min, info.min_allreduces, info.sum_allreduces) == (1, 0, 0, 0, 0)
    assert info.host_syncs == max(1, st.outer_iterations - 4)      # one synchronisation for the usual five iterations
    # an empty query set is the prior-only solve (cloud_matcher.cpp:153-154): the guess comes back
    pose0, st0 = m.align(q[:0], GUESS, with_stats=True)
    want0 = lo.CloudMatcher().align(g, q[:0], GUESS)
    assert np.array_equal(pose0, want0)
    m.close()


@pytest.mark.parametrize("sum_mode", ["fused", "nccl"])
def test_world_two_on_two_gpus(ctx, port, stream12, tmp_path, sum_mode):
    """fused: both exchanges through NVLink peer memory (atomicMin into the home rank inside the search + barrier + pull; the 32
    sums inside the LM step kernel); nccl: ncclAllReduce(MIN) and ncclAllReduce(SUM)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    kf, q = scene(port, stream12[0])
    np.savez(tmp_path / "scene.npz", kf=kf, q=q, guess=GUESS)
    env = dict(os.environ, LO_SHARDED_SUM=sum_mode, LO_SHARDED_MIN=sum_mode, LO_SHARDED_TIMEOUT_MS="20000")
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "sharded_nccl_worker.py"), str(r), "2", str(tmp_path)], env=env,
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=300)[0])
        except subprocess.TimeoutExpired:
            for pp in procs:
                pp.kill()
            raise
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    res = [json.load(open(tmp_path / f"result{r}.json")) for r in range(2)]
    g = lo.VoxelGrid(ctx, 0.2, 20)
    g.addCloud(kf)
    want, stats = lo.CloudMatcher().align(g, q, GUESS, with_stats=True)
    assert res[0]["voxels"] + res[1]["voxels"] == g.size() and res[0]["points"] + res[1]["points"] == g.numPoints()
    assert min(res[0]["voxels"], res[1]["voxels"]) > 0.25 * g.size()
    for run in range(3):
        a, b = res[0]["runs"][run], res[1]["runs"][run]
        assert a["pose_bits"] == b["pose_bits"] and a["cost"] == b["cost"]      # every rank holds the same result, bit for bit
        assert a["pose_bits"] == res[0]["runs"][0]["pose_bits"]                   # and repeated aligns repeat it
        assert a["outer"] == stats.outer_iterations and a["pairs"] == list(stats.pairs_log[:stats.outer_iterations])
        dt, dr = pose_error(np.array(a["pose"], np.float32), want)
        assert dt < 1e-6 and dr < 1e-6, (dt, dr)
    assert res[0]["empty_pose"] == res[1]["empty_pose"] == [float(v) for v in lo.CloudMatcher().align(g, q[:0], GUESS)]
    for r in res:
        assert r["info"]["world"] == 2 and r["info"]["fused_sum"] == (1 if sum_mode == "fused" else 0), r["info"]
        assert r["info"]["fused_min"] == (1 if sum_mode == "fused" else 0), r["info"]
        assert (r["info"]["min_allreduces"] if sum_mode == "nccl" else r["info"]["min_exchanges_fused"]) >= 15
        assert (r["info"]["sum_allreduces"] == 0) == (sum_mode == "fused") and (r["info"]["min_allreduces"] == 0) == (sum_mode == "fused")
