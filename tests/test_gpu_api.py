"""GPU parity for the entry points that close the VoxelGrid / utils / CloudClassifier interface:
setMaxPoints, getCorrespondence, pointTimeNormalize and transformNonRigid on their own, and the PAIR classify returns."""
import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from util import bits_equal, planar_scene, random_pn, sort_rows

pytestmark = pytest.mark.gpu
F = ("x", "y", "z", "nx", "ny", "nz")
IDENT = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)


def test_point_time_normalize_and_non_rigid_alone(ctx, port, stream12):
    scan = stream12[0][4]
    got, want = ctx.point_time_normalize(scan), port.time_normalize(scan)
    assert got.tobytes() == want.tobytes()                      # every field of every 32-byte record
    start = np.array([0.4, -0.02, 0.01, 0.99995, 0.001, -0.002, 0.009], np.float32)
    # transformNonRigid takes `time` as it is (cloud_transform.h:26-30): raw sweep times of ~0.4 s here, not [0, 1]
    for cloud in (scan, want):
        a, b = ctx.transform_non_rigid(cloud, start, IDENT), port.transform_nonrigid(cloud, start, IDENT)
        assert bits_equal(a, b, ("x", "y", "z", "time", "intensity")) and np.array_equal(a["ring"], b["ring"])


def test_classify_returns_the_pair(ctx, port, stream12):
    for k in (2, 7):
        deskewed = port.transform_nonrigid(port.time_normalize(stream12[0][k]), IDENT, IDENT)
        planar, un = ctx.classify_pair(deskewed)
        want_planar, want_un = port.classify_full(deskewed)
        assert len(planar) == len(want_planar) and bits_equal(planar, want_planar, F)
        # the second cloud: organised points that were looked at and did not become planar, intensity = curvature
        assert len(un) == len(want_un) and len(un) > 0
        assert bits_equal(un, want_un, ("x", "y", "z", "intensity", "time")) and np.array_equal(un["ring"], want_un["ring"])
        assert len(planar) + len(un) <= len(deskewed)


def test_classify_pair_ragged(ctx, port, stream12):
    scan = stream12[0][3]
    two = scan[(scan["ring"] == 3) | (scan["ring"] == 4)]
    planar, un = ctx.classify_pair(two)
    want_planar, want_un = port.classify_full(two)
    assert len(planar) == len(want_planar) and len(un) == len(want_un)
    assert bits_equal(un, want_un, ("x", "y", "z", "intensity"))
    p0, u0 = ctx.classify_pair(np.zeros(0, scan.dtype))
    assert len(p0) == 0 and len(u0) == 0


def test_get_correspondence_single_query(ctx, port, stream12):
    planar = planar_scene(port, stream12[0], 1)
    g, o = lo.VoxelGrid(ctx, 0.2, 20), port.grid(0.2, 20)
    g.addCloud(planar); o.add_cloud(planar)
    rng = np.random.default_rng(11)
    sel = rng.choice(len(planar), 300, replace=False)
    hits = 0
    for i, s in enumerate(sel):
        q = np.array([planar["x"][s], planar["y"][s], planar["z"][s]], np.float32) + rng.normal(0, 0.08, 3).astype(np.float32)
        for max_sq in (0.09, float(np.float32(0.3) * np.float32(0.3)), 1e-3, 0.0113371, 2.5):   # bounds that are not floats, too
            ok, origin, normal = g.getCorrespondence(q, max_sq)
            ok2, origin2, normal2 = o.get_correspondence(q, max_sq)
            assert ok == ok2, (i, max_sq)
            assert np.array_equal(origin, origin2) and np.array_equal(normal, normal2)
            hits += ok
    assert hits > 300
    far = np.array([1e4, 1e4, 1e4], np.float32)
    assert g.getCorrespondence(far, 0.09)[0] is False and o.get_correspondence(far, 0.09)[0] is False
    assert g.getCorrespondence(np.zeros(3, np.float32), -1.0)[0] is False


def test_set_max_points_only_changes_the_cap(ctx, port):
    """voxel_grid.h:56-59,82-91: voxels keep what they hold; later points are appended while a voxel holds fewer than the cap."""
    a, b, c = random_pn(4000, 21, spread=1.0), random_pn(4000, 22, spread=1.0), random_pn(4000, 23, spread=1.0)
    g, o = lo.VoxelGrid(ctx, 0.5, 4), port.grid(0.5, 4)
    for grid_call, orc_call in ((g.addCloud, o.add_cloud),):
        grid_call(a); orc_call(a)
    g.setMaxPoints(9); o.set_max_points(9)        # grow: the slabs are laid out again
    g.addCloud(b); o.add_cloud(b)
    assert g.size() == o.size() and g.numPoints() == o.num_points()
    assert np.array_equal(sort_rows(g.getCloud()), sort_rows(o.get_cloud()))
    g.setMaxPoints(2); o.set_max_points(2)        # shrink: nothing is dropped, fuller voxels just stop accepting
    g.addCloud(c); o.add_cloud(c)
    assert g.size() == o.size() and g.numPoints() == o.num_points()
    assert np.array_equal(sort_rows(g.getCloud()), sort_rows(o.get_cloud()))
    # correspondences still come out of the re-laid-out slabs identically
    q = np.array([a["x"][5], a["y"][5], a["z"][5]], np.float32)
    ok, origin, normal = g.getCorrespondence(q, 0.09)
    ok2, origin2, normal2 = o.get_correspondence(q, 0.09)
    assert ok and ok2 and np.array_equal(origin, origin2) and np.array_equal(normal, normal2)
    snap = g.snapshot()                           # a copy of a grid with a cap history restores
    g2 = lo.VoxelGrid(ctx, 0.5, 2)
    g2.restore(snap)
    assert g2.numPoints() == g.numPoints()


def test_staging_copy_helper_threads(ctx, stream12, monkeypatch):
    """lo_odometry_process on pageable scans: the staging copy with 0, 1 and 5 helper threads (LO_B200_COPY_THREADS is read when
    a handle's pool is created, i.e. at its first pageable scan) ends on the same poses; lo_odometry_host_threads reports
    caller + helpers."""
    poses = []
    for helpers in (0, 1, 5):
        monkeypatch.setenv("LO_B200_COPY_THREADS", str(helpers))
        od = lo.LidarOdometry(ctx)
        for s in stream12[0][:5]:
            od.processCloud(s)          # numpy memory: pageable
        assert od.hostThreads() == 1 + helpers
        poses.append(od.getCurrentPose().copy())
        del od
    assert np.array_equal(poses[0].view(np.uint32), poses[1].view(np.uint32)) and np.array_equal(poses[0].view(np.uint32), poses[2].view(np.uint32))
    monkeypatch.delenv("LO_B200_COPY_THREADS")
    assert 1 <= lo.LidarOdometry(ctx).hostThreads() <= 4
