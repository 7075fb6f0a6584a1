"""GPU: VoxelGrid::getCloud* in the iteration order of the reference's tsl::robin_map, position by position,
against the real container (oracle/_ref: the reference's voxel_grid.h over the vendored tsl::robin_map; the port, pinned to
it bit for bit by the CPU suite, where _ref is absent). Random key sets and adversarial ones: every key in the last few
buckets of the final table (one long cluster that wraps past the end), clusters forced at every growth step, many points per
voxel, several addCloud batches."""
import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
import oracle
from util import PN, planar_scene, sort_rows

pytestmark = pytest.mark.gpu
F = ("x", "y", "z", "nx", "ny", "nz")


@pytest.fixture(scope="module")
def real_map(port):
    return oracle.load("reference") if oracle.available("reference") else port


def hash22(ix, iy, iz):
    a = ix.astype(np.int64).astype(np.uint64) * np.uint64(73856093)
    b = iy.astype(np.int64).astype(np.uint64) * np.uint64(19349669)
    c = iz.astype(np.int64).astype(np.uint64) * np.uint64(83492791)
    return ((a ^ b ^ c) & np.uint64((1 << 22) - 1)).astype(np.int64)


def cloud_of(keys, rng, per_voxel=1):
    """One PointNormal per (key, repeat): voxel size 1, index i holds (i, i+1) for i >= 0 and (i-1, i) for i < 0."""
    k = np.repeat(keys, per_voxel, axis=0)
    frac = rng.uniform(0.1, 0.9, k.shape)
    p = np.where(k >= 0, k + frac, k - frac).astype(np.float32)
    out = np.zeros(len(p), PN)
    out["x"], out["y"], out["z"] = p.T
    out["nx"], out["ny"], out["nz"] = rng.normal(size=(3, len(p))).astype(np.float32)
    out["pad0"] = 1.0
    return out[rng.permutation(len(out))] if per_voxel > 1 else out


def check(ctx, orc, cloud, max_points=1, batches=1):
    g, o = lo.VoxelGrid(ctx, 1.0, max_points), orc.grid(1.0, max_points)
    for part in np.array_split(cloud, batches):
        g.addCloud(part)
        o.add_cloud(part)
    got, want = g.getCloud(reference_order=True), o.get_cloud()
    assert len(got) == len(want)
    for k in F:
        assert np.array_equal(got[k].view(np.uint32), want[k].view(np.uint32)), f"field {k}: {np.flatnonzero(got[k] != want[k])[:8]}"
    s_got, s_want = g.getSparseCloudWithoutNormals(reference_order=True), o.get_sparse_xyz()
    assert s_got.tobytes() == s_want.tobytes()


def candidate_keys(span=40):
    r = np.arange(-span, span)
    k = np.stack(np.meshgrid(r, r, r, indexing="ij"), -1).reshape(-1, 3)
    return k, hash22(k[:, 0], k[:, 1], k[:, 2])


def test_random_key_sets_every_size(ctx, real_map):
    rng = np.random.default_rng(1)
    keys, _ = candidate_keys(30)
    for n in (1, 2, 3, 4, 5, 7, 8, 9, 16, 17, 33, 100, 257, 1000, 4096, 4097, 14000):
        sel = keys[rng.choice(len(keys), n, replace=False)]
        check(ctx, real_map, cloud_of(sel, rng))


def test_forced_last_bucket_cluster_and_wrap(ctx, real_map):
    """All keys hash into the last `window` buckets of the final table: one cluster that runs past the end and wraps to the front,
    full of tie groups; inserted in random order, so groups are rotated by later insertions in front of them."""
    rng = np.random.default_rng(2)
    keys, h = candidate_keys(40)
    for n, window in ((6, 2), (40, 8), (200, 12), (200, 64), (1000, 30), (3000, 200)):
        C = 2
        while C < 2 * n:
            C <<= 1
        pool = np.flatnonzero((h & (C - 1)) >= C - window)
        assert len(pool) >= n
        sel = keys[rng.choice(pool, n, replace=False)]
        check(ctx, real_map, cloud_of(sel, rng))


def test_clusters_at_every_growth_step(ctx, real_map):
    """Keys whose low hash bits agree: they collide in every intermediate table (2, 4, 8, ... buckets) too, so each re-hash
    re-inserts long tie groups, some of them split by the wrap of the smaller table."""
    rng = np.random.default_rng(3)
    keys, h = candidate_keys(40)
    for n, low_bits, residues in ((64, 5, (31, 30)), (500, 6, (63, 0, 1)), (2000, 8, (255, 254, 7))):
        pool = np.flatnonzero(np.isin(h & ((1 << low_bits) - 1), residues))
        sel = keys[rng.choice(pool, n, replace=False)]
        check(ctx, real_map, cloud_of(sel, rng))


def test_many_points_per_voxel_and_batches(ctx, real_map):
    rng = np.random.default_rng(4)
    keys, _ = candidate_keys(12)
    sel = keys[rng.choice(len(keys), 1500, replace=False)]
    check(ctx, real_map, cloud_of(sel, rng, per_voxel=6), max_points=4, batches=5)
    check(ctx, real_map, cloud_of(sel[:300], rng, per_voxel=3), max_points=1, batches=3)


def test_scan_downsamplers(ctx, real_map, port, stream12):
    """The two uses of the order in lidar_odometry.cpp:37-50: 0.1 m and 0.3 m down-samplers of a scan's planar cloud."""
    for k in (1, 6):
        planar = planar_scene(port, stream12[0], k)
        for vs in (0.1, 0.3):
            g, o = lo.VoxelGrid(ctx, vs, 1), real_map.grid(vs, 1)
            g.addCloud(planar)
            o.add_cloud(planar)
            assert g.getCloud(reference_order=True).tobytes() == o.get_cloud().tobytes()
            assert g.getCloudWithoutNormals(reference_order=True).tobytes() == o.get_cloud_xyz().tobytes()


# ---- the block-local epoch replay (csrc/lo_robin_block.cuh: what the per-scan kernel runs) on its own ------------------------
def _real_order(real_map, keys, rng):
    """Iteration order of the real container for these voxel keys inserted in this order: index into `keys` per position."""
    cloud = cloud_of(keys, rng)
    o = real_map.grid(1.0, 1)
    o.add_cloud(cloud)
    got = o.get_cloud()
    lut = {(float(p["x"]), float(p["y"]), float(p["z"])): k for k, p in enumerate(cloud)}
    return np.array([lut[(float(p["x"]), float(p["y"]), float(p["z"]))] for p in got], np.uint32)


def _check_block_replay(ctx, real_map, keys, rng):
    h = hash22(keys[:, 0], keys[:, 1], keys[:, 2]).astype(np.uint32)
    got = ctx.robin_iteration_order(h)
    want = _real_order(real_map, keys, rng)
    assert np.array_equal(got, want), f"n = {len(keys)}: first difference at position {int(np.flatnonzero(got != want)[0])}"


def test_block_replay_random_sizes_across_every_epoch_boundary(ctx, real_map):
    rng = np.random.default_rng(21)
    keys, _ = candidate_keys(30)
    sizes = [1, 2, 3, 4, 5, 8, 9, 15, 16, 17, 31, 32, 33, 63, 64, 65, 127, 128, 129, 255, 256, 257, 511, 512, 513, 1000, 1024, 1025, 2048, 2049,
             4095, 4096, 4097, 8191, 8192, 8193, 12345, 14000, 16383, 16384]
    for n in sizes:
        _check_block_replay(ctx, real_map, keys[rng.choice(len(keys), n, replace=False)], rng)


def test_block_replay_forced_wrap_and_growth_collisions(ctx, real_map):
    rng = np.random.default_rng(22)
    keys, h = candidate_keys(40)
    for n, window in ((6, 2), (40, 8), (200, 12), (200, 64), (1000, 30), (1900, 200)):
        C = 2
        while C < 2 * n:
            C <<= 1
        pool = np.flatnonzero((h & (C - 1)) >= C - window)
        assert len(pool) >= n
        _check_block_replay(ctx, real_map, keys[rng.choice(pool, n, replace=False)], rng)
    # a degenerate key set (9 000 keys in 600 buckets: one run of 9 000 cells, beyond tsl's probe-distance limit) is refused,
    # not answered wrongly; the sequential replay behind VoxelGrid::getCloud* takes it (test_forced_last_bucket_cluster_and_wrap)
    pool = np.flatnonzero((h & 32767) >= 32768 - 600)
    sel = keys[rng.choice(pool, 9000, replace=False)]
    with pytest.raises(lo.LoError) as err:
        ctx.robin_iteration_order(hash22(sel[:, 0], sel[:, 1], sel[:, 2]).astype(np.uint32))
    assert err.value.code == lo.LO_E_RANGE
    for n, low_bits, residues in ((64, 5, (31, 30)), (500, 6, (63, 0, 1)), (2000, 8, (255, 254, 7)), (6000, 9, (511, 0, 1, 2, 300))):
        pool = np.flatnonzero(np.isin(h & ((1 << low_bits) - 1), residues))
        if len(pool) < n:
            continue
        _check_block_replay(ctx, real_map, keys[rng.choice(pool, n, replace=False)], rng)


def test_grid_replay_every_epoch_on_the_grid(ctx, real_map, monkeypatch):
    """csrc/lo_robin_grid.cuh with every epoch on the cooperative grid (LO_ROBIN_GRID=1), on sizes the block replay also takes:
    random sets across epoch boundaries, forced wrap, collisions at every growth step."""
    monkeypatch.setenv("LO_ROBIN_GRID", "1")
    rng = np.random.default_rng(31)
    keys, h = candidate_keys(30)
    for n in (1, 2, 3, 5, 16, 17, 33, 64, 65, 257, 1000, 1025, 4097, 9000, 16384):
        _check_block_replay(ctx, real_map, keys[rng.choice(len(keys), n, replace=False)], rng)
    keys, h = candidate_keys(40)
    for n, window in ((6, 2), (200, 12), (1000, 30), (1900, 200)):
        C = 2
        while C < 2 * n:
            C <<= 1
        pool = np.flatnonzero((h & (C - 1)) >= C - window)
        _check_block_replay(ctx, real_map, keys[rng.choice(pool, n, replace=False)], rng)
    for n, low_bits, residues in ((500, 6, (63, 0, 1)), (4000, 9, (511, 0, 1, 2, 300))):
        pool = np.flatnonzero(np.isin(h & ((1 << low_bits) - 1), residues))
        _check_block_replay(ctx, real_map, keys[rng.choice(pool, n, replace=False)], rng)


def test_grid_replay_large_key_sets(ctx, real_map):
    """Key sets beyond one block's shared memory (the update set of a 128-beam scan has ~180 k voxels): block 0 replays the
    first 2^14 elements, the grid the epochs after that. Position by position against the real container."""
    rng = np.random.default_rng(32)
    keys, h = candidate_keys(48)           # 884 736 candidate voxels
    for n in (16385, 20000, 32768, 32769, 70000, 180000, 300000):
        _check_block_replay(ctx, real_map, keys[rng.choice(len(keys), n, replace=False)], rng)
    # collisions at every growth step and a wrapping cluster at this size
    pool = np.flatnonzero(np.isin(h & 1023, (1023, 0, 1, 2, 3, 4, 500, 501)))
    _check_block_replay(ctx, real_map, keys[rng.choice(pool, 6500, replace=False)], rng)
    n = 40000
    pool = np.flatnonzero((h & 131071) >= 131072 - 60000)
    _check_block_replay(ctx, real_map, keys[rng.choice(pool, n, replace=False)], rng)


def test_block_replay_scan_update_sets(ctx, real_map, port, stream12):
    """The real use: the 0.1 m down-sampler of a scan's planar cloud (12-15 k voxels)."""
    for k in (1, 5, 9):
        planar = planar_scene(port, stream12[0], k)
        keys_all, h_all = ctx.voxel_keys(0.1, np.ascontiguousarray(planar[["x", "y", "z", "pad0"]]).view(lo.XYZ) if False else _xyz_of(planar))
        _, first = np.unique(keys_all, axis=0, return_index=True)
        first.sort()                                  # voxels in order of first arrival = insertion order of the reference
        got = ctx.robin_iteration_order(h_all[first])
        o = real_map.grid(0.1, 1)
        o.add_cloud(planar)
        want = o.get_cloud()
        sel = planar[first[got]]
        assert len(sel) == len(want) and all(np.array_equal(sel[f].view(np.uint32), want[f].view(np.uint32)) for f in ("x", "y", "z"))


def _xyz_of(pn):
    out = np.zeros(len(pn), lo.XYZ)
    out["x"], out["y"], out["z"], out["pad"] = pn["x"], pn["y"], pn["z"], 1.0
    return out


def _run_against_real_container(ctx, real_map, scans):
    od, oc = lo.LidarOdometry(ctx), real_map.odometry()
    orders = []
    for k, s in enumerate(scans):
        od.processCloud(s)
        oc.process(s)
        st = od.stats()
        orders.append(st.update_order)
        a, b = od.getKeyFrameCloudWithNormals(), oc.keyframe_pn()
        assert len(a) == len(b), k
        assert np.array_equal(sort_rows(a), sort_rows(b)), f"key frames differ after scan {k}"
    return orders


def test_per_scan_kernel_update_order(ctx, real_map, stream12):
    """The fused per-scan kernel (csrc/lo_pipeline.cu) feeds keyframe_.addCloud(update) (lidar_odometry.cpp:42,70) in the
    counting-sort order where that provably gives the reference's key frame, and in the replayed order otherwise: after every
    scan the key frame holds the same points as the pipeline over the real container, and no scan reports an approximation."""
    orders = _run_against_real_container(ctx, real_map, stream12[0])
    assert all(o in (lo.LO_ORDER_SORTED, lo.LO_ORDER_REPLAYED) for o in orders), orders


def test_pair_check_finds_the_tie_group_pair(ctx, real_map):
    """Scan 8 of the 120-scan stream (tools/list_replayed.py) has a group of equal ideal bucket with two update points in one
    key-frame voxel: the pair check of the locate pass must find it and hand the insert to the replay kernel; the other scans
    stay on the counting sort; the key frame equals the real-container pipeline's after every scan."""
    from lidar_odometry_demo_b200 import synthetic as S
    n = 120                                              # the scene's length follows the stream's length: build that stream's head
    lidar = S.LidarModel.vlp16()
    pos, q, _ = S.trajectory(n + 1, 99)
    scene = S.Scene(1234, length=max(60.0, pos[-1, 0] + 140.0))
    scans = [S.generate_scan(scene, lidar, pos[k], q[k], pos[k + 1], q[k + 1], 1234 + k, t_offset=0.1 * k) for k in range(10)]
    orders = _run_against_real_container(ctx, real_map, scans)
    assert orders[8] == lo.LO_ORDER_REPLAYED, orders
    assert all(o == lo.LO_ORDER_SORTED for k, o in enumerate(orders) if k != 8), orders


def test_per_scan_kernel_replay_forced():
    """The same with the block replay forced on every scan (LO_B200_REPLAY_ALWAYS is read when the first scan is queued, so
    this test owns a fresh library state only if it runs in its own process: the switch is latched per process)."""
    import subprocess
    import sys
    import os
    code = (
        "import sys, numpy as np\n"
        "sys.path.insert(0, 'tests')\n"
        "import lidar_odometry_demo_b200 as lo, oracle\n"
        "from lidar_odometry_demo_b200 import synthetic as S\n"
        "from util import sort_rows\n"
        "orc = oracle.load('reference' if oracle.available('reference') else 'port')\n"
        "scans, _, _ = S.vlp16_stream(8)\n"
        "ctx = lo.Context(0)\n"
        "od, oc = lo.LidarOdometry(ctx), orc.odometry()\n"
        "for k, s in enumerate(scans):\n"
        "    od.processCloud(s); oc.process(s)\n"
        "    assert od.stats().update_order == lo.LO_ORDER_REPLAYED, (k, od.stats().update_order)\n"
        "    a, b = od.getKeyFrameCloudWithNormals(), oc.keyframe_pn()\n"
        "    assert len(a) == len(b) and np.array_equal(sort_rows(a), sort_rows(b)), k\n"
        "    d = np.abs(od.getCurrentPose().astype(np.float64) - oc.pose().astype(np.float64)).max()\n"
        "    assert d < 1e-5, (k, d)\n"
        "print('replayed', len(scans))\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, LO_B200_REPLAY_ALWAYS="1"), cwd=root, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "replayed 8" in r.stdout


def test_large_scan_update_order_replayed_on_the_grid():
    """A 64-beam scan (131 072 rays, update voxels of 0.05 m: update sets of several 10^4 points, beyond one block's shared
    memory) through scan_kernel_large with the replay forced on every scan: the grid-wide replay (csrc/lo_robin_grid.cuh)
    must leave the key frame exactly as the pipeline over the real container does."""
    import os
    import subprocess
    import sys
    code = (
        "import sys, ctypes, numpy as np\n"
        "sys.path.insert(0, 'tests')\n"
        "import lidar_odometry_demo_b200 as lo, oracle\n"
        "from lidar_odometry_demo_b200 import synthetic as S\n"
        "from util import sort_rows\n"
        "orc = oracle.load('reference' if oracle.available('reference') else 'port')\n"
        "lidar = S.LidarModel(64, 2048, -25.0, 15.0)\n"
        "pos, q, _ = S.trajectory(4, 99)\n"
        "scene = S.Scene(4321, length=400.0, scale=2.0)\n"
        "scans = [S.generate_scan(scene, lidar, pos[k], q[k], pos[k + 1], q[k + 1], 4321 + k, t_offset=0.1 * k) for k in range(3)]\n"
        "p = lo.Params(); lo.load_library().lo_params_default(ctypes.byref(p))\n"
        "p.keyframe_voxel_size, p.keyframe_update_voxel_size, p.keyframe_matching_voxel_size = 0.1, 0.05, 0.05\n"
        "op = orc.default_params()\n"
        "op.keyframe_voxel_size, op.keyframe_update_voxel_size, op.keyframe_matching_voxel_size = 0.1, 0.05, 0.05\n"
        "ctx = lo.Context(0)\n"
        "od, oc = lo.LidarOdometry(ctx, p, keyframe_voxels=1 << 20, scan_points=1 << 17), orc.odometry(op)\n"
        "for k, s in enumerate(scans):\n"
        "    od.processCloud(s); oc.process(s)\n"
        "    st = od.stats()\n"
        "    assert st.n_update > 16384, st.n_update\n"
        "    assert st.update_order == lo.LO_ORDER_REPLAYED, (k, st.update_order)\n"
        "    a, b = od.getKeyFrameCloudWithNormals(), oc.keyframe_pn()\n"
        "    assert len(a) == len(b) and np.array_equal(sort_rows(a), sort_rows(b)), k\n"
        "    d = np.abs(od.getCurrentPose().astype(np.float64) - oc.pose().astype(np.float64)).max()\n"
        "    assert d < 1e-5, (k, d)\n"
        "print('replayed', len(scans))\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, LO_B200_REPLAY_ALWAYS="1"), cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "replayed 3" in r.stdout
