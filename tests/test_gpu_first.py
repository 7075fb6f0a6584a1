"""First GPU parity tests: voxel keys, grid insert, correspondences, normal equations, align — CUDA through
the C ABI against the CPU oracle on the same inputs."""
import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from util import PN, XYZ, bits_equal, planar_scene, pn_to_xyz, pose_error, random_pn, sort_rows

pytestmark = pytest.mark.gpu
F = ("x", "y", "z", "nx", "ny", "nz")


def test_voxel_keys_bit_exact(ctx, port):
    rng = np.random.default_rng(1)
    pts = np.concatenate([(rng.normal(size=(50000, 3)) * 30).astype(np.float32),
                          np.array([[-0.05, -0.25, 0.0], [0.2, 0.4, -0.2], [1e-8, -1e-8, 99.99999]], np.float32)])
    cloud = np.zeros(len(pts), XYZ)
    cloud["x"], cloud["y"], cloud["z"] = pts.T
    for vs in (0.1, 0.2, 0.3, 0.5):
        keys, h22 = ctx.voxel_keys(vs, cloud)
        ref_keys = port.get_indices(vs, pts)
        assert np.array_equal(keys, ref_keys)
        sel = slice(0, 2000)
        assert [port.hash(*k) for k in ref_keys[sel]] == list(h22[sel])


@pytest.mark.parametrize("vs,cap", [(0.5, 1), (1.0, 3), (2.0, 20)])
def test_grid_insert_matches_oracle(ctx, port, vs, cap):
    pn = random_pn(20000, 2)
    g = lo.VoxelGrid(ctx, vs, cap)
    o = port.grid(vs, cap)
    g.addCloud(pn[:12000]); o.add_cloud(pn[:12000])
    g.addCloud(pn[12000:]); o.add_cloud(pn[12000:])
    assert g.size() == o.size() and g.numPoints() == o.num_points()
    assert np.array_equal(sort_rows(g.getCloud()), sort_rows(o.get_cloud()))
    # radius cleanup + re-insert (exercises tombstones and payload reuse)
    c = np.array([1, 2, 3], np.float32)
    g.radiusCleanup(c, 25.0); o.radius_cleanup(c, 25.0)
    assert g.size() == o.size() and g.numPoints() == o.num_points()
    g.addCloud(pn[::-1]); o.add_cloud(pn[::-1])
    assert g.size() == o.size() and g.numPoints() == o.num_points()
    assert np.array_equal(sort_rows(g.getCloud()), sort_rows(o.get_cloud()))
    assert np.array_equal(sort_rows(g.getSparseCloudWithoutNormals(), ("x", "y", "z")),
                          sort_rows(o.get_sparse_xyz(), ("x", "y", "z")))


def test_correspondences_bit_exact(ctx, port, stream12):
    scans = stream12[0]
    kf_cloud = planar_scene(port, scans, 0)
    q_cloud = pn_to_xyz(planar_scene(port, scans, 1))[::3]
    g = lo.VoxelGrid(ctx, 0.2, 20); o = port.grid(0.2, 20)
    g.addCloud(kf_cloud); o.add_cloud(kf_cloud)
    pose = np.array([0.03, -0.01, 0.02, 0.99998, 0.001, -0.002, 0.005], np.float32)
    pg, vg = g.findMatchingPairs(q_cloud, pose, 0.3)
    po, vo = o.find_matching_pairs(q_cloud, pose, 0.3)
    assert vo.sum() > 1000
    assert np.array_equal(vg, vo)
    assert np.array_equal(pg.view(np.uint64), po.view(np.uint64))  # source, origin and normal: identical bits


def test_normal_equations(ctx, port, stream12):
    scans = stream12[0]
    kf_cloud = planar_scene(port, scans, 0)
    q_cloud = pn_to_xyz(planar_scene(port, scans, 1))[::3]
    g = lo.VoxelGrid(ctx, 0.2, 20); o = port.grid(0.2, 20)
    g.addCloud(kf_cloud); o.add_cloud(kf_cloud)
    pose = np.array([0.03, -0.01, 0.02, 0.99998, 0.001, -0.002, 0.005], np.float32)
    x = np.array([pose[3], pose[4], pose[5], pose[6], pose[0], pose[1], pose[2]], np.float64)
    prior = np.array([0.0, 0.0, 0.0])
    cost, grad, jtj, nv = g.nnResidual(q_cloud, pose, x, prior)
    po, vo = o.find_matching_pairs(q_cloud, pose, 0.3)
    c0, g0, h0 = port.evaluate(po[vo], x, prior)
    assert nv == int(vo.sum())
    assert abs(cost - c0) <= 1e-9 * abs(c0)            # north_star: residuals within 1e-5 relative
    assert np.allclose(grad, g0, rtol=1e-9, atol=1e-9 * np.abs(g0).max())
    assert np.allclose(jtj, h0, rtol=1e-9, atol=1e-9 * np.abs(h0).max())


def test_align_matches_oracle(ctx, port, stream12):
    scans = stream12[0]
    kf_cloud = planar_scene(port, scans, 0)
    g = lo.VoxelGrid(ctx, 0.2, 20); o = port.grid(0.2, 20)
    g.addCloud(kf_cloud); o.add_cloud(kf_cloud)
    ds = port.grid(0.3, 1)
    ds.add_cloud(planar_scene(port, scans, 1))
    q_cloud = ds.get_cloud_xyz()
    for guess in ([0, 0, 0, 1, 0, 0, 0], [0.1, 0.05, -0.02, 0.9999619, 0, 0, 0.0087265]):
        guess = np.array(guess, np.float32)
        pose_g, stats = lo.CloudMatcher().align(g, q_cloud, guess, with_stats=True)
        pose_o, log = o.align(q_cloud, guess)
        dt, dr = pose_error(pose_g, pose_o)
        assert stats.outer_iterations == len(log)
        assert dt < 1e-4 and dr < 1e-4                 # north_star: 1e-4 m / 1e-4 rad per scan
        assert [int(v) for v in log[:, 0]] == list(stats.pairs_log[:len(log)])
