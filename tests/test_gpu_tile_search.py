"""The tile search (csrc/lo_tile_search.cuh: own voxel first, neighbours pruned by the best distance so far) must give
exactly what the exhaustive 27-voxel visit of the reference gives (src/voxel_grid.h:164-204): same correspondences,
hence the same normal equations and the same align result — on a map of planar patches with a large query set
(BASELINE.json config 3/4 shape, scaled so the CPU oracle finishes in seconds)."""
import os

import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from lidar_odometry_demo_b200 import synthetic as S
from util import pose_error

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def patch_map():
    big = S.planar_patch_map(600_000, seed=777, extent=(120.0, 120.0, 20.0))
    rng = np.random.default_rng(11)
    sel = rng.choice(len(big), 48_000, replace=False)
    q = np.zeros(len(sel), S.XYZ)
    # most queries near stored points, some shifted into empty space (no correspondence) and some onto voxel borders
    q["x"], q["y"], q["z"] = big["x"][sel] + 0.03, big["y"][sel] - 0.02, big["z"][sel] + 0.01
    far = rng.random(len(q)) < 0.15
    q["z"][far] += rng.uniform(0.25, 0.6, int(far.sum())).astype(np.float32)
    edge = rng.random(len(q)) < 0.1
    q["x"][edge] = (np.round(q["x"][edge] / 0.2) * 0.2).astype(np.float32)
    return big, q


def test_tile_normal_equations_equal_exhaustive(ctx, patch_map, monkeypatch):
    big, q = patch_map
    g = lo.VoxelGrid(ctx, 0.2, 20, capacity_hint=len(big) // 8)
    g.addCloud(big)
    pose = np.array([0.02, -0.01, 0.015, 0.9999875, 0.0, 0.0, 0.005], np.float32)
    x = np.array([pose[3], pose[4], pose[5], pose[6], pose[0], pose[1], pose[2]], np.float64)
    monkeypatch.delenv("LO_B200_NN_EXHAUSTIVE", raising=False)
    cost_t, grad_t, hess_t, nv_t = g.nnResidual(q, pose, x, x[4:])
    visited_t = g.last_candidates
    monkeypatch.setenv("LO_B200_NN_EXHAUSTIVE", "1")
    cost_e, grad_e, hess_e, nv_e = g.nnResidual(q, pose, x, x[4:])
    visited_e = g.last_candidates
    assert nv_t == nv_e and 0 < nv_t < len(q)            # same set of valid pairs (some queries have none)
    assert visited_t < visited_e                           # the pruning did prune
    # same pairs, different order of the double sums
    assert abs(cost_t - cost_e) <= 1e-12 * abs(cost_e)
    assert np.allclose(grad_t, grad_e, rtol=1e-10, atol=1e-9)
    assert np.allclose(hess_t, hess_e, rtol=1e-10, atol=1e-9)


def test_large_align_matches_oracle(ctx, port, patch_map):
    """m >= 8 queries per warp: align_kernel searches with the tile search."""
    big, q = patch_map
    g = lo.VoxelGrid(ctx, 0.2, 20, capacity_hint=len(big) // 8)
    o = port.grid(0.2, 20)
    g.addCloud(big)
    o.add_cloud(big)
    guess = np.array([0.03, -0.02, 0.01, 0.9999875, 0.0, 0.0, 0.005], np.float32)
    pose_g, stats = lo.CloudMatcher().align(g, q, guess, with_stats=True)
    pose_o, log = o.align(q, guess)
    dt, dr = pose_error(pose_g, pose_o)
    assert stats.outer_iterations == len(log)
    assert [int(v) for v in log[:, 0]] == list(stats.pairs_log[:len(log)])   # same number of pairs in every outer iteration
    assert dt < 1e-4 and dr < 1e-4                                          # north_star: 1e-4 m / 1e-4 rad
