"""CPU suite (-m "not gpu"): pins the oracle port against (a) the known-answer vectors of the reference's own
tests (/root/reference/test/test.cpp, cited per test), (b) the golden fixtures under tests/golden/ generated
from the reference's own headers by tools/make_golden.py, and (c) oracle/_ref itself where it is built."""
import hashlib
import os

import numpy as np
import pytest

from lidar_odometry_demo_b200 import synthetic as S
from util import PN, XYZ, bits_equal, pn_to_xyz, pose_error, random_pn, sort_rows

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
IDENT = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
F = ("x", "y", "z", "nx", "ny", "nz")


def gold(name):
    return np.load(os.path.join(GOLD, name))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).view(np.uint8).tobytes()).hexdigest()


def pn_from6(a):
    out = np.zeros(len(a), PN)
    for c, k in enumerate(F):
        out[k] = a[:, c]
    return out


def aa(angle, axis):
    """Eigen::Quaternionf(AngleAxisf(angle, axis.normalized())) as pose7 rotation part."""
    axis = np.asarray(axis, np.float64)
    axis = axis / np.linalg.norm(axis)
    return np.concatenate([[np.cos(angle / 2)], np.sin(angle / 2) * axis]).astype(np.float32)


def mat4(p):
    """Independent double-precision isometry of a pose7 (what the reference's tests build with Eigen::Isometry3f)."""
    t, (w, x, y, z) = np.asarray(p[:3], np.float64), np.asarray(p[3:], np.float64)
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    M = np.eye(4)
    M[:3, :3], M[:3, 3] = R, t
    return M


def same_pose(p7, M, tol_t):
    Mp = mat4(p7)
    assert np.linalg.norm(Mp[:3, 3] - M[:3, 3]) < tol_t
    assert np.abs(Mp[:3, :3] - M[:3, :3]).max() < 1e-5   # ASSERT_FLOAT_EQ(|q.q'|, 1) equivalent


# ---- the reference's own unit tests, replayed on the oracle --------------------------------------------
@pytest.mark.parametrize("kind", ["port", "reference"])
def test_voxelgrid_unique_points(kind, request):
    """test/test.cpp:26-55"""
    orc = request.getfixturevalue("port" if kind == "port" else "ref")
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, 0, 0], [0, -1, 0], [0, 0, -1]], np.float32)
    cloud = np.zeros(7, PN)
    cloud["x"], cloud["y"], cloud["z"] = pts.T
    g = orc.grid(0.5, 1)
    g.add_cloud(cloud)
    assert g.size() == 7
    out = g.get_cloud()
    assert len(out) == 7
    assert np.array_equal(sort_rows(out, ("x", "y", "z")), sort_rows(cloud, ("x", "y", "z")))


@pytest.mark.parametrize("kind", ["port", "reference"])
def test_voxelgrid_duplicate_points(kind, request):
    """test/test.cpp:57-75"""
    orc = request.getfixturevalue("port" if kind == "port" else "ref")
    cloud = np.zeros(4, PN)
    cloud["x"] = [0, 1, 0, 1]
    g = orc.grid(0.5, 1)
    g.add_cloud(cloud)
    assert g.size() == 2
    out = g.get_cloud()
    assert len(out) == 2 and out["x"][0] != out["x"][1]


POSE_PAIRS = [
    (np.r_[0, 0, 0, 1, 0, 0, 0], np.r_[0, 0, 0, 1, 0, 0, 0]),
    (np.r_[0, 0, 0, aa(0.2, (0, 0, 1))], np.r_[0, 0, 0, aa(0.2, (0, 0, 1))]),
    (np.r_[0, 0, 0, aa(0, (0, 0, 1))], np.r_[1, 0, 0, aa(np.pi * 0.5, (0, 0, 1))]),
    (np.r_[1, 0, 0, aa(0, (0, 0, 1))], np.r_[1, 1, 1, aa(-np.pi, (0, 0, 1))]),
    (np.r_[100, 100, 100, aa(0, (0, 0, 1))], np.r_[150, 150, 150, aa(0, (0, 0, 1))]),
    (np.r_[100, 100, 100, aa(0.1, (0, 0, 1))], np.r_[150, 150, 150, aa(-0.2, (0, 0, 1))]),
    (np.r_[1, 0.5, -0.5, aa(0.456, (0.1, 0.2, 1))], np.r_[-1, -0.6, 0, aa(-0.245, (-0.2, 0, 0))]),
]


@pytest.mark.parametrize("kind", ["port", "reference"])
def test_pose3d_compose_relative_inverse(kind, request):
    """test/test.cpp:77-149 with the same tolerances (compose 1e-6, relative / inverse 1e-4)."""
    orc = request.getfixturevalue("port" if kind == "port" else "ref")
    for a, b in POSE_PAIRS:
        a, b = a.astype(np.float32), b.astype(np.float32)
        Ma, Mb = mat4(a), mat4(b)
        same_pose(orc.pose_compose(a, b), Ma @ Mb, 1e-6 * max(1.0, np.abs(a[:3]).max()))
        same_pose(orc.pose_relative_to(a, b), np.linalg.inv(Ma) @ Mb, 1e-4)
        same_pose(orc.pose_inverse(a), np.linalg.inv(Ma), 1e-4)
        same_pose(orc.pose_inverse(b), np.linalg.inv(Mb), 1e-4)


@pytest.mark.parametrize("kind", ["port", "reference"])
def test_cloud_transformer_rigid(kind, request):
    """test/test.cpp:151-189: 7 points x 6 poses, per-coordinate tolerance 1e-7 against an independent transform."""
    orc = request.getfixturevalue("port" if kind == "port" else "ref")
    pts = np.array([[0, 0, 0], [1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], np.float32)
    cloud = S.to_xyz(pts)
    d = np.pi / 180.0
    cases = [np.r_[0, 0, 0, aa(0, (0, 0, 1))], np.r_[0, 0, 0, aa(80 * d, (0, 0, 1))], np.r_[0, 0, 0, aa(80 * d, (0, 1, 0))],
             np.r_[0, 0, 0, aa(90 * d, (1, 0, 1))], np.r_[1, 1, 1, aa(45 * d, (0, 0.5, 0.5))], np.r_[-2, 2, 0, aa(45 * d, (0.5, 0.5, 0))]]
    for p in cases:
        p = p.astype(np.float32)
        got = orc.transform_xyz(cloud, p)
        M = mat4(p)
        want = pts.astype(np.float64) @ M[:3, :3].T + M[:3, 3]
        g3 = np.stack([got["x"], got["y"], got["z"]], 1)
        assert np.abs(g3 - want).max() < 3e-7   # 1e-7 in the reference, vs PCL's float path; ours vs a double path


def test_hash_known_answers(port):
    """SURVEY F10 probes of IndicesHash (src/voxel_grid.h:31-38)."""
    assert port.hash(-1, 0, 0) == 1641379
    assert port.hash(1, 2, 3) == 363058
    # truncation toward zero of an IEEE float divide (src/voxel_grid.h:68-75)
    k = port.get_indices(0.2, np.array([[-0.05, -0.25, 0.0], [0.2, 0.4, -0.2]], np.float32))
    assert k[0].tolist() == [0, -1, 0]
    assert k[1, 0] == 1


# ---- golden fixtures generated from the reference's own headers (tools/make_golden.py) ------------------
def test_golden_keys_and_hashes(port):
    v = gold("vectors.npz")
    for vs in (0.1, 0.2, 0.3, 0.5):
        keys = port.get_indices(vs, v["key_points"])
        assert np.array_equal(keys, v[f"keys_{vs}"])
        assert np.array_equal(np.array([port.hash(*k) for k in keys], np.uint32), v[f"hash_{vs}"])


def test_golden_robin_order(port):
    """getCloud() of a VoxelGrid(vs,1) walks tsl::robin_map's bucket array."""
    v = gold("vectors.npz")
    p1 = pn_from6(v["planar1"])
    for vs in (0.1, 0.3):
        g = port.grid(vs, 1)
        g.add_cloud(p1)
        got = g.get_cloud()
        want = p1[v[f"robin_order_{vs}"]]
        assert len(got) == len(want) and bits_equal(got, want, F)


def test_golden_correspondences_and_normal_equations(port):
    v = gold("vectors.npz")
    p0, p1 = pn_from6(v["planar0"]), pn_from6(v["planar1"])
    kf = port.grid(0.2, 20)
    kf.add_cloud(p0)
    q = pn_to_xyz(p1[v["query_index"]])
    pairs9, valid = kf.find_matching_pairs(q, v["pose"], 0.3)
    assert np.array_equal(valid, v["valid"])
    want = p0[v["match_index"]]
    assert np.array_equal(pairs9[valid, 3:6], np.stack([want["x"], want["y"], want["z"]], 1).astype(np.float64))
    assert np.array_equal(pairs9[valid, 6:9], np.stack([want["nx"], want["ny"], want["nz"]], 1).astype(np.float64))
    pose = v["pose"]
    x = np.array([pose[3], pose[4], pose[5], pose[6], pose[0], pose[1], pose[2]], np.float64)
    cost, grad, jtj = port.evaluate(pairs9[valid], x, np.zeros(3))
    assert abs(cost - float(v["cost"])) <= 1e-12 * abs(cost)
    assert np.allclose(grad, v["grad"], rtol=1e-11, atol=1e-11 * np.abs(v["grad"]).max())
    assert np.allclose(jtj, v["jtj"], rtol=1e-11, atol=1e-11 * np.abs(v["jtj"]).max())


def test_golden_align(port):
    v = gold("vectors.npz")
    p0, p1 = pn_from6(v["planar0"]), pn_from6(v["planar1"])
    kf = port.grid(0.2, 20)
    kf.add_cloud(p0)
    q = pn_to_xyz(p1[v["query_index"]])
    pose, log = kf.align(q, v["align_guess"])
    assert len(log) == len(v["align_log"])
    assert np.array_equal(log[:, 0], v["align_log"][:, 0])       # pairs per outer iteration
    assert np.array_equal(pose.view(np.uint32), v["align_pose"].view(np.uint32))


def test_golden_odometry(port):
    """LidarOdometry::processCloud over 12 synthetic scans, port vs the reference's headers: identical poses."""
    o = gold("odometry.npz")
    scans, _, _ = S.vlp16_stream(12)
    if [sha(s) for s in scans] != list(o["scan_sha"]):
        pytest.skip("synthetic generator produced different bits on this host (numpy/libm version)")
    od = port.odometry()
    for k, s in enumerate(scans):
        counts, _ = od.process(s)
        assert counts.astype(np.int64).tolist() == o["counts"][k].tolist()
        assert np.array_equal(od.pose().view(np.uint32), o["poses"][k].view(np.uint32)), k
    assert len(od.keyframe(False)) == int(o["keyframe_voxels"])


def test_golden_config1(port):
    """BASELINE config 1: the MatchingTest flow (test/test.cpp:191-264) on intersection00056.pcd."""
    c = gold("config1.npz")
    xyz = S.to_xyz(gold("pcd_intersection00056.npz")["xyz"])
    nrm = port.pca_normals(xyz, 0.25)
    ok = ~np.isnan(nrm).any(1)
    assert int(ok.sum()) == int(c["n_valid_normals"]) and sha(nrm[ok]) == str(c["normals_sha"])
    pn = np.zeros(int(ok.sum()), PN)
    for i, k in enumerate("xyz"):
        pn[k], pn["n" + k] = xyz[k][ok], nrm[ok, i]
    kf = port.grid(0.25, 20)
    kf.add_cloud(pn)
    assert kf.size() == int(c["kf_voxels"]) and kf.num_points() == int(c["kf_points"])
    ds = port.grid(0.5, 1)
    ds.add_cloud_xyz(xyz)
    sub = ds.get_cloud_xyz()
    assert bits_equal(sub, xyz[c["sub_index"]], ("x", "y", "z"))
    for g, want, n_outer in zip(c["guesses"], c["final_poses"], c["outer_iterations"]):
        cloud = port.transform_xyz(sub, port.pose_inverse(g))
        pose, log = kf.align(cloud, IDENT)
        assert len(log) == int(n_outer)
        assert np.array_equal(pose.view(np.uint32), want.view(np.uint32))
        # the reference's own acceptance bound (test.cpp:254-262): final.relativeTo(guess)
        rel = port.pose_relative_to(pose, g)
        assert np.linalg.norm(rel[:3]) < 0.05
        assert 1.0 - abs(float(np.dot(pose[3:], g[3:]))) < 0.01


# ---- port vs oracle/_ref on randomised inputs ------------------------------------------------------------
def test_port_vs_ref_grid_and_order(port, ref):
    pn = random_pn(30000, 7, spread=15.0)
    for vs, cap in ((0.5, 1), (0.3, 1), (1.0, 5), (2.0, 20)):
        a, b = port.grid(vs, cap), ref.grid(vs, cap)
        a.add_cloud(pn[:20000]); b.add_cloud(pn[:20000])
        a.add_cloud(pn[20000:]); b.add_cloud(pn[20000:])
        assert a.size() == b.size() and a.num_points() == b.num_points()
        assert bits_equal(a.get_cloud(), b.get_cloud(), F)            # same robin iteration order
        c = np.array([1, -2, 0.5], np.float32)
        a.radius_cleanup(c, 20.0); b.radius_cleanup(c, 20.0)
        assert a.size() == b.size()
        assert bits_equal(a.get_cloud(), b.get_cloud(), F)            # backward-shift erase keeps order identical
        a.add_cloud(pn[::-1][:5000]); b.add_cloud(pn[::-1][:5000])
        assert bits_equal(a.get_cloud(), b.get_cloud(), F)
        assert bits_equal(a.get_sparse_xyz(), b.get_sparse_xyz(), ("x", "y", "z"))


def test_port_vs_ref_point_stages(port, ref, stream12):
    scan = stream12[0][4]
    start = np.array([0.4, -0.02, 0.01, 0.99995, 0.001, -0.002, 0.009], np.float32)
    a, b = port.time_normalize(scan), ref.time_normalize(scan)
    assert a.tobytes() == b.tobytes()
    da, db = port.transform_nonrigid(a, start, IDENT), ref.transform_nonrigid(b, start, IDENT)
    assert da.tobytes() == db.tobytes()
    (pa, ua), (pb, ub) = port.classify(da), ref.classify(db)
    assert ua == ub and len(pa) == len(pb) and bits_equal(pa, pb, F)
    fa, fb = port.range_filter(pa, 4.0, 80.0), ref.range_filter(pb, 4.0, 80.0)
    assert len(fa) == len(fb) and bits_equal(fa, fb, F)
    ta, tb = port.transform_with_normals(fa, start), ref.transform_with_normals(fb, start)
    assert bits_equal(ta, tb, F)


def test_empty_and_degenerate_inputs(port):
    g = port.grid(0.2, 20)
    assert g.size() == 0 and len(g.get_cloud()) == 0
    q = S.to_xyz(np.zeros((5, 3), np.float32))
    pairs, valid = g.find_matching_pairs(q, IDENT, 0.3)
    assert not valid.any()
    pose, log = g.align(q, np.array([1, 2, 3, 1, 0, 0, 0], np.float32))   # prior-only solve stays at the guess
    assert np.allclose(pose, [1, 2, 3, 1, 0, 0, 0], atol=1e-6) and len(log) >= 5
    # max_points = 0 still stores the first point of a new voxel (src/voxel_grid.h:82-91)
    z = port.grid(0.5, 0)
    z.add_cloud(random_pn(100, 1))
    assert z.size() == z.num_points()
