"""GPU parity against the committed golden fixtures (tests/golden/, generated from the reference's own headers by
tools/make_golden.py) — these run on the GPU box, where /root/reference does not exist."""
import hashlib
import os

import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from lidar_odometry_demo_b200 import synthetic as S
from util import PN, XYZ, bits_equal, pn_to_xyz, pose_error

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
IDENT = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
F = ("x", "y", "z", "nx", "ny", "nz")


def gold(name):
    return np.load(os.path.join(GOLD, name))


def pn_from6(a):
    out = np.zeros(len(a), PN)
    for c, k in enumerate(F):
        out[k] = a[:, c]
    return out


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).view(np.uint8).tobytes()).hexdigest()


def test_keys_and_hashes(ctx):
    v = gold("vectors.npz")
    cloud = S.to_xyz(v["key_points"])
    for vs in (0.1, 0.2, 0.3, 0.5):
        keys, h22 = ctx.voxel_keys(vs, cloud)
        assert np.array_equal(keys, v[f"keys_{vs}"]) and np.array_equal(h22, v[f"hash_{vs}"])   # bit-exact


def test_correspondences_normal_equations_align(ctx):
    v = gold("vectors.npz")
    p0, p1 = pn_from6(v["planar0"]), pn_from6(v["planar1"])
    kf = lo.VoxelGrid(ctx, 0.2, 20)
    kf.addCloud(p0)
    q = pn_to_xyz(p1[v["query_index"]])
    pairs9, valid = kf.findMatchingPairs(q, v["pose"], 0.3)
    assert np.array_equal(valid, v["valid"])                                            # correspondence set: exact
    want = p0[v["match_index"]]
    assert np.array_equal(pairs9[valid, 3:6], np.stack([want["x"], want["y"], want["z"]], 1).astype(np.float64))
    assert np.array_equal(pairs9[valid, 6:9], np.stack([want["nx"], want["ny"], want["nz"]], 1).astype(np.float64))
    pose = v["pose"]
    x = np.array([pose[3], pose[4], pose[5], pose[6], pose[0], pose[1], pose[2]], np.float64)
    cost, grad, jtj, nv = kf.nnResidual(q, pose, x, np.zeros(3))
    assert nv == int(valid.sum())
    assert abs(cost - float(v["cost"])) <= 1e-9 * abs(cost)                              # north_star: 1e-5 relative
    assert np.allclose(grad, v["grad"], rtol=1e-9, atol=1e-9 * np.abs(v["grad"]).max())
    assert np.allclose(jtj, v["jtj"], rtol=1e-9, atol=1e-9 * np.abs(v["jtj"]).max())
    out, stats = lo.CloudMatcher().align(kf, q, v["align_guess"], with_stats=True)
    dt, dr = pose_error(out, v["align_pose"])
    assert stats.outer_iterations == len(v["align_log"])
    assert list(stats.pairs_log[:stats.outer_iterations]) == [int(k) for k in v["align_log"][:, 0]]
    assert dt < 1e-6 and dr < 1e-6                                                      # north_star: 1e-4 m / 1e-4 rad


def test_odometry_poses(ctx):
    o = gold("odometry.npz")
    scans, _, _ = S.vlp16_stream(12)
    if [sha(s) for s in scans] != list(o["scan_sha"]):
        pytest.skip("synthetic generator produced different bits on this host")
    od = lo.LidarOdometry(ctx)
    for k, s in enumerate(scans):
        od.processCloud(s)
        st = od.stats()
        assert [st.n_points, st.n_planar, st.n_update, st.n_query] == o["counts"][k].tolist()
        dt, dr = pose_error(od.getCurrentPose(), o["poses"][k])
        assert dt < 1e-6 and dr < 1e-6, (k, dt, dr)
    assert od.stats().keyframe_voxels == int(o["keyframe_voxels"])


def test_config1_matching_test_flow(ctx, port):
    """BASELINE config 1: test/test.cpp:191-264 on intersection00056.pcd; K7 plane fit on the GPU."""
    c = gold("config1.npz")
    xyz = S.to_xyz(gold("pcd_intersection00056.npz")["xyz"])
    nrm = ctx.pca_normals(xyz, 0.25)
    want = port.pca_normals(xyz, 0.25)
    ok = ~np.isnan(want).any(1)
    assert np.array_equal(~np.isnan(nrm).any(1), ok) and int(ok.sum()) == int(c["n_valid_normals"])
    # same neighbourhoods, covariance in double, analytic eigenvector: agreement to 1e-5 except where the two smallest
    # eigenvalues nearly coincide (normal direction ill-conditioned)
    dots = np.abs(np.sum(nrm[ok].astype(np.float64) * want[ok], axis=1))
    assert np.mean(dots > 1 - 1e-5) > 0.995
    pn = np.zeros(int(ok.sum()), PN)
    for i, k in enumerate("xyz"):
        pn[k], pn["n" + k] = xyz[k][ok], want[ok, i]        # oracle normals: the align comparison below isolates K4-K6
    kf = lo.VoxelGrid(ctx, 0.25, 20)
    kf.addCloud(pn)
    assert kf.size() == int(c["kf_voxels"]) and kf.numPoints() == int(c["kf_points"])
    ds = lo.VoxelGrid(ctx, 0.5, 1)
    ds.addCloudWithoutNormals(xyz)
    sub = xyz[c["sub_index"]]        # the reference's robin_map order of the sub-sampled cloud
    assert len(ds.getCloudWithoutNormals()) == len(sub)
    for g, fin, n_outer in zip(c["guesses"], c["final_poses"], c["outer_iterations"]):
        cloud = ctx.transform_xyz(sub, port.pose_inverse(g))
        assert bits_equal(cloud, port.transform_xyz(sub, port.pose_inverse(g)), ("x", "y", "z"))
        out, stats = lo.CloudMatcher().align(kf, cloud, IDENT, with_stats=True)
        dt, dr = pose_error(out, fin)
        assert stats.outer_iterations == int(n_outer)
        assert dt < 1e-5 and dr < 1e-5, (g, dt, dr)        # north_star: 1e-4 m / 1e-4 rad
