"""Edge cases of LidarOdometry::processCloud through the per-scan kernel (and its stage-by-stage twin) against the CPU
oracle: empty scans, scans with no usable point, ragged ring sets, non-finite returns, the size where the two
pipelines hand over, empty query sets for align / the NN kernel."""
import os
import subprocess
import sys

import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from lidar_odometry_demo_b200 import synthetic as S
from util import pose_error

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_both(ctx, port, scans):
    od, oc = lo.LidarOdometry(ctx), port.odometry()
    worst = (0.0, 0.0)
    for k, sc in enumerate(scans):
        od.processCloud(sc)
        counts, _ = oc.process(sc)
        dt, dr = pose_error(od.getCurrentPose(), oc.pose())
        worst = (max(worst[0], dt), max(worst[1], dr))
        st = od.stats()
        assert (st.n_planar, st.n_update, st.n_query) == (counts[1], counts[2], counts[3]), f"scan {k}: planar/update/query counts differ"
    return worst, od, oc


def test_empty_and_unusable_scans(ctx, port, stream12):
    scans = stream12[0]
    empty = np.zeros(0, scans[0].dtype)
    near = scans[2].copy()                      # every return inside lidar_min_range: nothing survives the range filter
    for k in "xyz":
        near[k] *= np.float32(0.01)
    seq = [empty, scans[0], empty, scans[1], near, scans[3], scans[4]]
    (dt, dr), od, oc = run_both(ctx, port, seq)
    assert dt < 1e-4 and dr < 1e-4


def test_ragged_rings_and_non_finite_returns(ctx, port, stream12):
    scans = stream12[0]
    two = scans[2][(scans[2]["ring"] == 3) | (scans[2]["ring"] == 7)]
    holes = scans[3].copy()
    holes["x"][::97] = np.nan                   # dropped returns as NaN / inf, as real drivers deliver them
    holes["y"][5::131] = np.inf
    few = scans[4][:40]                         # fewer points than a ring has columns
    (dt, dr), od, oc = run_both(ctx, port, [scans[0], scans[1], two, holes, few, scans[5]])
    assert dt < 1e-4 and dr < 1e-4


def test_pipeline_handover_size(ctx, port, stream12):
    """Scans up to 524 288 points run in the per-scan kernel, larger ones stage by stage: same results on both sides. (The
    scan is a VLP16 revolution repeated: 16 rings of ~34 k columns each.)"""
    scans = stream12[0]
    base = np.tile(scans[0], 19)
    limit = 1 << 19
    assert len(base) > limit + 1
    a, b = base[:limit].copy(), base[:limit + 1].copy()
    l0 = ctx.kernel_launches
    (dt, dr), od, oc = run_both(ctx, port, [a, b, a])
    assert dt < 1e-4 and dr < 1e-4
    assert 10 < ctx.kernel_launches - l0 < 200        # two scans = one launch each (+ table growth), one scan = many


def test_stage_by_stage_pipeline_matches_oracle():
    """LO_B200_STAGED=1 (read once per process) routes every scan through the stage-by-stage kernels."""
    code = """
import numpy as np, oracle, lidar_odometry_demo_b200 as lo
from lidar_odometry_demo_b200 import synthetic as S
import sys; sys.path.insert(0, 'tests')
from util import pose_error
scans, _, _ = S.vlp16_stream(6)
ctx = lo.Context(0); od = lo.LidarOdometry(ctx); oc = oracle.load('port').odometry()
for sc in scans:
    od.processCloud(sc); oc.process(sc)
    dt, dr = pose_error(od.getCurrentPose(), oc.pose())
    assert dt < 1e-4 and dr < 1e-4, (dt, dr)
print('ok', ctx.kernel_launches)
"""
    env = dict(os.environ, LO_B200_STAGED="1")
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert int(r.stdout.split()[-1]) > 6 * 10          # many launches per scan: it really took the staged path


def test_empty_query_sets(ctx, port, stream12):
    from util import planar_scene
    kf = planar_scene(port, stream12[0], 0)
    g = lo.VoxelGrid(ctx, 0.2, 20)
    g.addCloud(kf)
    none = np.zeros(0, S.XYZ)
    guess = np.array([0.1, 0.0, 0.0, 1, 0, 0, 0], np.float32)
    pose, st = lo.CloudMatcher().align(g, none, guess, with_stats=True)
    o = port.grid(0.2, 20)
    o.add_cloud(kf)
    pose_o, log = o.align(none, guess)
    dt, dr = pose_error(pose, pose_o)
    assert dt < 1e-6 and dr < 1e-6 and st.valid_pairs == 0 and st.outer_iterations == len(log)
    x = np.array([1, 0, 0, 0, 0.1, 0, 0], np.float64)
    cost, grad, hess, nv = g.nnResidual(none, guess, x, x[4:])
    assert nv == 0 and cost == 0.0
    empty_grid = lo.VoxelGrid(ctx, 0.2, 20)
    q = np.zeros(100, S.XYZ)
    pose, st = lo.CloudMatcher().align(empty_grid, q, guess, with_stats=True)     # prior-only solve (cloud_matcher.cpp:153)
    assert st.valid_pairs == 0 and np.allclose(pose[:3], guess[:3], atol=1e-6)


def test_process_next_equals_process(ctx, stream12):
    """lo_odometry_process_next (next scan uploaded behind the current one) gives the poses of lo_odometry_process, also
    when the announced next scan is not the one that arrives."""
    scans = [np.ascontiguousarray(s) for s in stream12[0][:8]]
    a, b = lo.LidarOdometry(ctx), lo.LidarOdometry(ctx)
    for k, sc in enumerate(scans):
        a.processCloud(sc)
        nxt = scans[k + 1] if k + 1 < len(scans) else None
        if k == 3:
            nxt = scans[0]          # wrong announcement: must be ignored at the next call
        b.processCloudWithNext(sc.ctypes.data, len(sc), nxt.ctypes.data if nxt is not None else None, len(nxt) if nxt is not None else 0)
        assert np.array_equal(a.getCurrentPose(), b.getCurrentPose()), k


def test_submit_wait_pipeline_equals_process(ctx, stream12):
    """lo_odometry_submit / lo_odometry_wait with two scans in flight, then with the queue filled to its depth: the poses,
    counts and key frame of the one-call lo_odometry_process, scan by scan (an empty scan first, so the bootstrap decision
    is taken on the device twice)."""
    scans = [np.zeros(0, stream12[0][0].dtype)] + [np.ascontiguousarray(s) for s in stream12[0][:9]]
    a, b = lo.LidarOdometry(ctx), lo.LidarOdometry(ctx)
    ref = []
    for sc in scans:
        a.processCloud(sc)
        st = a.stats()
        ref.append((a.getCurrentPose().copy(), st.n_planar, st.n_update, st.n_query, st.outer_iterations, st.keyframe_voxels))
    got = []

    def collect():
        b.wait()
        st = b.stats()
        got.append((b.getCurrentPose().copy(), st.n_planar, st.n_update, st.n_query, st.outer_iterations, st.keyframe_voxels))

    b.submit(scans[0].ctypes.data, len(scans[0]))
    for k in range(1, len(scans)):
        b.submit(scans[k].ctypes.data, len(scans[k]))            # scan k queued behind scan k-1
        if k == 3:
            with pytest.raises(lo.LoError):                        # a synchronous call is refused while scans are in flight
                b.processCloud(scans[k])
        collect()
    collect()
    assert len(got) == len(ref)
    # the whole queue: QUEUE_DEPTH scans submitted before the first wait, one more is refused and nothing is lost
    c = lo.LidarOdometry(ctx)
    c.processCloud(scans[0])
    c.processCloud(scans[1])
    deep = []
    assert lo.QUEUE_DEPTH == 8 and len(scans) >= 2 + lo.QUEUE_DEPTH
    for k in range(2, 2 + lo.QUEUE_DEPTH):
        c.submit(scans[k].ctypes.data, len(scans[k]))
    with pytest.raises(lo.LoError) as e:
        c.submit(scans[2].ctypes.data, len(scans[2]))
    assert e.value.code == lo.LO_E_BUSY
    for k in range(2, 2 + lo.QUEUE_DEPTH):
        c.wait()
        st = c.stats()
        deep.append((c.getCurrentPose().copy(), st.n_planar, st.n_update, st.n_query, st.outer_iterations, st.keyframe_voxels))
    for k, (r, g) in enumerate(zip(ref[2:2 + lo.QUEUE_DEPTH], deep)):
        assert np.array_equal(r[0], g[0]) and r[1:] == g[1:], (k, r, g)
    for k, (r, g) in enumerate(zip(ref, got)):
        assert np.array_equal(r[0], g[0]) and r[1:] == g[1:], (k, r, g)
    with pytest.raises(lo.LoError):
        b.wait()                                                   # nothing in flight
    b.processCloud(scans[1])                                       # the synchronous call works again


def test_config3_scan_size(ctx, port):
    """BASELINE.json config 3: 128 beams x 2048 columns (~245 k returns per scan), voxels 0.1 / 0.05 / 0.05 — the largest scans
    of the contract (per-scan kernel with the tile search in align; bucket bins of the robin order span several chunks): counts
    and poses as the oracle."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from bench_c3 import c3_params, c3_stream
    scans = c3_stream(3)
    assert len(scans[0]) > 200_000
    od = lo.LidarOdometry(ctx, c3_params(lo))
    op = port.default_params()
    op.keyframe_voxel_size, op.keyframe_update_voxel_size, op.keyframe_matching_voxel_size = 0.1, 0.05, 0.05
    oc = port.odometry(op)
    for k, sc in enumerate(scans):
        od.processCloud(sc)
        counts, _ = oc.process(sc)
        st = od.stats()
        assert (st.n_planar, st.n_update, st.n_query) == (counts[1], counts[2], counts[3]), k
        dt, dr = pose_error(od.getCurrentPose(), oc.pose())
        assert dt < 1e-4 and dr < 1e-4, (k, dt, dr)
