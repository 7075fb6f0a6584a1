#!/usr/bin/env python
"""Generates tests/golden/*.npz from the reference's own headers (oracle/_ref/liblo_ref.so, built from
/root/reference by `make -C oracle ref`). Run in the build container only; the fixtures are committed so the
GPU box (which has no /root/reference) can check both the oracle port and the CUDA path against them.

    python tools/make_golden.py

Fixtures
  pcd_intersection00056.npz  the xyz columns of the reference's test/test_data/intersection00056.pcd (config 1 input)
  config1.npz                MatchingTest flow (test/test.cpp:191-264) on that cloud: sizes, 7 guesses, final poses, logs
  vectors.npz                voxel keys / hashes, robin_map iteration order of the two down-samplers, correspondences,
                             normal equations and one align on a synthetic VLP16 pair (config 2 shape)
  odometry.npz               LidarOdometry::processCloud over 12 synthetic VLP16 scans: pose and N/P/U/M per scan
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from lidar_odometry_demo_b200 import synthetic as S  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
PCD = "/root/reference/test/test_data/intersection00056.pcd"
IDENT = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)


def quat_axis_angle(angle, axis):
    """Eigen::Quaternionf(AngleAxisf(angle, axis)) in float."""
    axis = np.asarray(axis, np.float32)
    h = np.float32(0.5) * np.float32(angle)
    s, c = np.float32(np.sin(h)), np.float32(np.cos(h))
    return np.array([c, s * axis[0], s * axis[1], s * axis[2]], np.float32)


def matching_test_guesses():
    """The 7 guess poses of test/test.cpp:235-243 (translation, AngleAxis about z)."""
    rows = [((0, 0, 0), 0.0), ((0, 0, 0.1), 0.0), ((0.1, 0.1, 0.1), 0.0), ((-0.1, -0.1, -0.1), 0.0),
            ((0.1, -0.1, 0), 0.0), ((0, 0, 0), -1.0), ((-0.2, 0, 0), 2.0)]
    out = []
    for t, deg in rows:
        q = quat_axis_angle(np.float32(deg * np.pi / 180.0), (0, 0, 1)) if deg else np.array([1, 0, 0, 0], np.float32)
        out.append(np.concatenate([np.asarray(t, np.float32), q]))
    return np.stack(out).astype(np.float32)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).view(np.uint8).tobytes()).hexdigest()


def index_of(rows, table):
    """position in `table` of every row of `rows` (exact float bits on x,y,z)."""
    key = lambda c: (c["x"].view(np.uint32).astype(np.uint64) << 40) ^ (c["y"].view(np.uint32).astype(np.uint64) << 20) ^ c["z"].view(np.uint32)
    kt, kr = key(table), key(rows)
    order = np.argsort(kt, kind="stable")
    pos = np.searchsorted(kt[order], kr)
    idx = order[np.clip(pos, 0, len(order) - 1)]
    assert np.array_equal(kt[idx], kr)
    return idx.astype(np.int32)


def main():
    os.makedirs(OUT, exist_ok=True)
    ref, port = oracle.load("reference"), oracle.load("port")
    assert ref.kind == "reference"

    # ---- config 1 ---------------------------------------------------------------------------------
    xyz3 = S.read_pcd_xyz(PCD)
    np.savez_compressed(os.path.join(OUT, "pcd_intersection00056.npz"), xyz=xyz3)
    xyz = S.to_xyz(xyz3)
    nrm = port.pca_normals(xyz, 0.25)           # PCL is absent: restated, parity unpinned
    ok = ~np.isnan(nrm).any(1)
    pn = np.zeros(int(ok.sum()), S.PN)
    for c, k in enumerate("xyz"):
        pn[k], pn["n" + k] = xyz[k][ok], nrm[ok, c]
    kf = ref.grid(0.25, 20)
    kf.add_cloud(pn)
    ds = ref.grid(0.5, 1)
    ds.add_cloud_xyz(xyz)
    sub = ds.get_cloud_xyz()
    guesses = matching_test_guesses()
    finals, outers, pairs = [], [], []
    for g in guesses:
        cloud = ref.transform_xyz(sub, ref.pose_inverse(g))
        pose, log = kf.align(cloud, IDENT)
        finals.append(pose); outers.append(len(log)); pairs.append(np.pad(log[:, 0], (0, 35 - len(log))))
    np.savez_compressed(os.path.join(OUT, "config1.npz"), n_valid_normals=int(ok.sum()), normals_sha=sha(nrm[ok]),
                        normals_sample=nrm[::64], kf_voxels=kf.size(), kf_points=kf.num_points(), sub_index=index_of(sub, xyz),
                        guesses=guesses, final_poses=np.stack(finals), outer_iterations=np.array(outers),
                        pairs_log=np.stack(pairs).astype(np.int32))

    # ---- synthetic VLP16 pair ---------------------------------------------------------------------
    scans, _, _ = S.vlp16_stream(12)

    def planar_of(k):
        d = ref.transform_nonrigid(ref.time_normalize(scans[k]), IDENT, IDENT)
        planar, _ = ref.classify(d)
        return ref.range_filter(planar, 4.0, 80.0)

    p0, p1 = planar_of(0), planar_of(1)
    rng = np.random.default_rng(1)
    pts = np.concatenate([(rng.normal(size=(4096, 3)) * 30).astype(np.float32),
                          np.array([[-0.05, -0.25, 0.0], [0.2, 0.4, -0.2], [1e-8, -1e-8, 99.99999], [-0.2, 0.6, 0.3]], np.float32)])
    vec = {"key_points": pts, "scan_sha": np.array([sha(s) for s in scans[:2]]),
           "planar0": np.stack([p0[k] for k in ("x", "y", "z", "nx", "ny", "nz")], 1),
           "planar1": np.stack([p1[k] for k in ("x", "y", "z", "nx", "ny", "nz")], 1)}
    for vs in (0.1, 0.2, 0.3, 0.5):
        keys = ref.get_indices(vs, pts)
        vec[f"keys_{vs}"] = keys
        vec[f"hash_{vs}"] = np.array([ref.hash(*k) for k in keys], np.uint32)
    for vs in (0.1, 0.3):
        g = ref.grid(vs, 1)
        g.add_cloud(p1)
        vec[f"robin_order_{vs}"] = index_of(g.get_cloud(), p1)
    kf = ref.grid(0.2, 20)
    kf.add_cloud(p0)
    g = ref.grid(0.3, 1)
    g.add_cloud(p1)
    q = g.get_cloud_xyz()
    pose = np.array([0.03, -0.01, 0.02, 0.99998, 0.001, -0.002, 0.005], np.float32)
    pairs9, valid = kf.find_matching_pairs(q, pose, 0.3)
    x = np.array([pose[3], pose[4], pose[5], pose[6], pose[0], pose[1], pose[2]], np.float64)
    cost, grad, jtj = ref.evaluate(pairs9[valid], x, np.zeros(3))
    guess = np.array([0.1, 0.05, -0.02, 0.9999619, 0, 0, 0.0087265], np.float32)
    apose, alog = kf.align(q, guess)
    vec.update(query_index=index_of(q, p1), pose=pose, valid=valid, match_index=index_of(
        S.to_xyz(pairs9[valid, 3:6].astype(np.float32)), p0), cost=cost, grad=grad, jtj=jtj, align_guess=guess,
        align_pose=apose, align_log=alog)
    np.savez_compressed(os.path.join(OUT, "vectors.npz"), **vec)

    # ---- odometry over 12 scans -------------------------------------------------------------------
    od = ref.odometry()
    poses, counts = [], []
    for s in scans:
        c, _ = od.process(s)
        poses.append(od.pose()); counts.append(c)
    np.savez_compressed(os.path.join(OUT, "odometry.npz"), scan_sha=np.array([sha(s) for s in scans]), poses=np.stack(poses),
                        counts=np.stack(counts).astype(np.int64), keyframe_voxels=len(od.keyframe(False)))
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
