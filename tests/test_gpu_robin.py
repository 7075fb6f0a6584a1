"""GPU: VoxelGrid::getCloud* in the iteration order of the reference's tsl::robin_map, position by position,
against the real container (oracle/_ref: the reference's voxel_grid.h over the vendored tsl::robin_map; the port, pinned to
it bit for bit by the CPU suite, where _ref is absent). Random key sets and adversarial ones: every key in the last few
buckets of the final table (one long cluster that wraps past the end), clusters forced at every growth step, many points per
voxel, several addCloud batches."""
import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
import oracle
from util import PN, planar_scene

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
