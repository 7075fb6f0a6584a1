"""GPU: the headless driver (lidar_odometry_demo_b200/apps/lo_replay.cpp) — recorded scans from
.bin / .pcd files through the C++ facade, trajectory and key-frame files out, checkpoint and resume. It must give the
poses the library gives when called directly."""
import json
import os
import subprocess

import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from lidar_odometry_demo_b200 import build as lib_build

pytestmark = pytest.mark.gpu


def write_pcd_xyzirt(path, scan):
    head = (f"# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z _ intensity ring _ time _\nSIZE 4 4 4 4 4 2 2 4 4\n"
            f"TYPE F F F U F U U F U\nCOUNT 1 1 1 1 1 1 1 1 1\nWIDTH {len(scan)}\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS {len(scan)}\nDATA binary\n")
    with open(path, "wb") as f:
        f.write(head.encode())
        f.write(scan.tobytes())


def test_replay_matches_direct_calls(ctx, stream12, tmp_path):
    exe = lib_build.build_apps()
    scans = stream12[0][:10]
    od = lo.LidarOdometry(ctx)
    want = []
    for sc in scans:
        od.processCloud(sc)
        want.append(od.getCurrentPose().copy())
    want = np.array(want)
    files = []
    for k, sc in enumerate(scans):
        path = tmp_path / (f"scan{k:03d}.pcd" if k % 2 else f"scan{k:03d}.bin")
        if k % 2:
            write_pcd_xyzirt(path, sc)
        else:
            sc.tofile(path)
        files.append(str(path))
    (tmp_path / "list.txt").write_text("# ten scans\n" + "\n".join(files) + "\n")

    def run(*args):
        r = subprocess.run([exe, *args], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        return json.loads(r.stdout.strip().splitlines()[-1])

    for mode in ([], ["--queue"]):
        out = run("--out", str(tmp_path / "tum.txt"), "--keyframe-pcd", str(tmp_path / "kf.pcd"), "--full-keyframe-pcd", str(tmp_path / "kff.pcd"),
                  "--deskewed-pcd", str(tmp_path / "desk.pcd"), "--list", str(tmp_path / "list.txt"), *mode)
        tum = np.loadtxt(tmp_path / "tum.txt")
        assert tum.shape == (10, 8) and out["scans"] == 10 and out["points"] == sum(len(s) for s in scans)
        got = np.concatenate([tum[:, 1:4], tum[:, 7:8], tum[:, 4:7]], axis=1).astype(np.float32)   # -> tx ty tz qw qx qy qz
        assert np.array_equal(got, want), mode                                                      # %.9g round-trips float32
        assert np.allclose(tum[:, 0], 0.1 * np.arange(10))
        assert np.array_equal(np.array(out["final_pose"], np.float32), want[-1])
        for name, n in (("kf.pcd", len(od.getKeyFrameCloud())), ("kff.pcd", len(od.getFullKeyFrameCloud())), ("desk.pcd", len(scans[-1]))):
            head = open(tmp_path / name, "rb").read(400).decode("latin1")
            assert f"POINTS {n}\n" in head, name
    # KITTI output + checkpoint after 6 scans, resume for the other 4
    run("--format", "kitti", "--out", str(tmp_path / "k1.txt"), "--checkpoint-out", str(tmp_path / "state.bin"), *files[:6])
    out = run("--format", "kitti", "--out", str(tmp_path / "k2.txt"), "--checkpoint-in", str(tmp_path / "state.bin"), *files[6:])
    assert np.array_equal(np.array(out["final_pose"], np.float32), want[-1])
    k = np.vstack([np.loadtxt(tmp_path / "k1.txt"), np.loadtxt(tmp_path / "k2.txt")]).reshape(10, 3, 4)
    assert np.array_equal(k[:, :, 3].astype(np.float32), want[:, :3])
    # params.yaml path and a bad file
    y = tmp_path / "params.yaml"
    y.write_text("/**:\n  ros__parameters:\n    lidar_odometry:\n      lidar_min_range: 4.0\n      lidar_max_range: 80.0\n      keyframe_voxel_size: 0.2\n"
                 "      keyframe_max_points_cnt: 20\n      keyframe_matching_voxel_size: 0.3\n      keyframe_update_voxel_size: 0.1\n"
                 "      keyframe_cleanup_range: 80.0\n      angular_divergence_threshold: 5.0\n")
    out = run("--params", str(y), *files)
    assert np.array_equal(np.array(out["final_pose"], np.float32), want[-1])
    (tmp_path / "odd.bin").write_bytes(b"123")
    r = subprocess.run([exe, str(tmp_path / "odd.bin")], capture_output=True, text=True)
    assert r.returncode == 1 and "32-byte" in r.stderr
