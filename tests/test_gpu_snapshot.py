"""Snapshot / restore of the device table and of a running odometry. The reference has no
checkpoint (VoxelGrid and LidarOdometry are plain values, src/lidar_odometry.h:82-85), so the contract is the one a
copy of those values would give: a restored object answers every later call exactly like the original."""
import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from util import pn_to_xyz, random_pn, sort_rows

pytestmark = pytest.mark.gpu


def test_grid_snapshot_round_trip(ctx):
    a = lo.VoxelGrid(ctx, 0.5, 6)
    a.addCloud(random_pn(40000, 1))
    a.radiusCleanup([0.0, 0.0, 0.0], 30.0)          # tombstones + a free list in the snapshot
    a.addCloud(random_pn(20000, 2, spread=12.0))
    blob = a.snapshot()
    b = lo.VoxelGrid(ctx, 0.5, 6, capacity_hint=64)  # smaller table: restore re-sizes it
    b.restore(blob)
    assert (a.size(), a.numPoints()) == (b.size(), b.numPoints())
    assert np.array_equal(sort_rows(a.getCloud()), sort_rows(b.getCloud()))
    q = pn_to_xyz(random_pn(5000, 3, spread=12.0))
    ident = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
    pa, va = a.findMatchingPairs(q, ident, 0.4)
    pb, vb = b.findMatchingPairs(q, ident, 0.4)
    assert va.sum() > 100 and np.array_equal(va, vb) and pa.tobytes() == pb.tobytes()
    # both keep evolving the same way
    more = random_pn(15000, 4, spread=15.0)
    a.addCloud(more)
    b.addCloud(more)
    a.radiusCleanup([1.0, 2.0, 0.0], 20.0)
    b.radiusCleanup([1.0, 2.0, 0.0], 20.0)
    assert (a.size(), a.numPoints()) == (b.size(), b.numPoints())
    assert np.array_equal(sort_rows(a.getCloud()), sort_rows(b.getCloud()))


def test_grid_restore_rejects_foreign_blobs(ctx):
    a = lo.VoxelGrid(ctx, 0.5, 6)
    a.addCloud(random_pn(1000, 5))
    blob = a.snapshot()
    with pytest.raises(lo.LoError) as e:
        lo.VoxelGrid(ctx, 0.5, 7).restore(blob)      # different max_points
    assert e.value.code == lo.LO_E_ARG
    bad = blob.copy()
    bad[0] ^= 0xFF
    for broken in (bad, blob[:-4], blob[:40]):
        with pytest.raises(lo.LoError) as e:
            a.restore(broken)
        assert e.value.code == lo.LO_E_IO
    a.restore(blob)                                   # the grid is still usable
    assert a.numPoints() > 0


def test_odometry_checkpoint_continues_bit_identically(ctx, stream12):
    scans = stream12[0]
    a = lo.LidarOdometry(ctx)
    for sc in scans[:6]:
        a.processCloud(sc)
    blob = a.snapshot()
    tail_a = []
    for sc in scans[6:12]:
        a.processCloud(sc)
        st = a.stats()
        tail_a.append((a.getCurrentPose().tobytes(), st.n_planar, st.n_update, st.n_query, st.outer_iterations, st.keyframe_voxels))
    b = lo.LidarOdometry(ctx)
    b.processCloud(scans[0])                          # restore replaces whatever state the object had
    b.restore(blob)
    assert b.getTempCloud().size == 0
    tail_b = []
    for sc in scans[6:12]:
        b.processCloud(sc)
        st = b.stats()
        tail_b.append((b.getCurrentPose().tobytes(), st.n_planar, st.n_update, st.n_query, st.outer_iterations, st.keyframe_voxels))
    assert tail_a == tail_b
    assert np.array_equal(sort_rows(a.getKeyFrameCloudWithNormals()), sort_rows(b.getKeyFrameCloudWithNormals()))
    p = lo.Params()
    ctx.lib.lo_params_default(lo.C.byref(p))
    p.keyframe_voxel_size = 0.25
    with pytest.raises(lo.LoError) as e:
        lo.LidarOdometry(ctx, p).restore(blob)
    assert e.value.code == lo.LO_E_ARG
