"""lidar_odometry_demo_b200 — B200-native scan-to-keyframe registration.

Python is only a test / bench driver here: this module binds the C ABI of liblo_b200.so
(include/lo_b200.h) with ctypes and mirrors the reference's class names (VoxelGrid, CloudMatcher,
LidarOdometry; /root/reference/src/*.h) so that the parity tests read like the reference's own tests.
There is no CPU fallback: importing works anywhere (so the symbol table can be checked without a GPU),
but creating a Context without a CUDA device raises.
"""
import ctypes as C
import os

import numpy as np

from .synthetic import PN, XYZ, XYZIRT  # noqa: F401  (wire dtypes)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LO_B200_LIB") or os.path.join(_HERE, "liblo_b200.so")   # the override: A/B runs of another build

QUEUE_DEPTH = 8   # LO_ODOMETRY_QUEUE_DEPTH
LO_OK, LO_E_CUDA, LO_E_ARG, LO_E_CAPACITY, LO_E_RANGE, LO_E_IO, LO_E_BUSY = 0, -1, -2, -3, -4, -5, -6


class LoError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"liblo_b200 error {code}: {text}")
        self.code = code


class Pair(C.Structure):
    _fields_ = [("source_point_local", C.c_double * 3), ("plane_origin", C.c_double * 3), ("plane_normal", C.c_double * 3)]


class NormalEq(C.Structure):
    _fields_ = [("cost", C.c_double), ("g", C.c_double * 6), ("jtj", C.c_double * 36), ("n_valid", C.c_uint64),
                ("candidates", C.c_uint64), ("occupied", C.c_uint64)]


class AlignStats(C.Structure):
    _fields_ = [("outer_iterations", C.c_int32), ("lm_iterations", C.c_int32), ("valid_pairs", C.c_int32),
                ("termination", C.c_int32), ("final_cost", C.c_double), ("last_step_norm", C.c_double),
                ("pose_log", (C.c_float * 7) * 35), ("pairs_log", C.c_int32 * 35)]


class Params(C.Structure):
    _fields_ = [("lidar_min_range", C.c_float), ("lidar_max_range", C.c_float), ("keyframe_voxel_size", C.c_float),
                ("keyframe_max_points_cnt", C.c_uint64), ("keyframe_matching_voxel_size", C.c_float),
                ("keyframe_update_voxel_size", C.c_float), ("keyframe_cleanup_range", C.c_float),
                ("angular_divergence_threshold", C.c_float)]


class ShardedInfo(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("fused_sum", C.c_int32), ("nccl_version", C.c_int32),
                ("min_allreduces", C.c_uint64), ("sum_allreduces", C.c_uint64), ("host_syncs", C.c_uint64), ("evaluations", C.c_uint64),
                ("fused_min", C.c_int32), ("reserved", C.c_int32), ("min_exchanges_fused", C.c_uint64)]


class KeyframeInfo(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("voxels", "points", "tombstones", "capacity_slots", "capacity_voxels", "table_rebuilds",
                                         "payload_growths")]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


LO_ORDER_SORTED, LO_ORDER_REPLAYED, LO_ORDER_APPROXIMATE = 0, 1, 2   # lo_scan_stats.update_order


class ScanStats(C.Structure):
    _fields_ = [("n_points", C.c_uint32), ("n_planar", C.c_uint32), ("n_update", C.c_uint32), ("n_query", C.c_uint32),
                ("outer_iterations", C.c_int32), ("unstable", C.c_int32), ("keyframe_voxels", C.c_uint32),
                ("update_order", C.c_uint32)]


_vp, _sz, _fl, _i = C.c_void_p, C.c_size_t, C.c_float, C.c_int
# every symbol include/lo_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "lo_ctx_create": (_i, [_i, C.POINTER(_vp)]), "lo_ctx_destroy": (None, [_vp]), "lo_last_error": (C.c_char_p, []),
    "lo_ctx_stream": (_vp, [_vp]), "lo_ctx_synchronize": (_i, [_vp]), "lo_ctx_kernel_launches": (C.c_uint64, [_vp]),
    "lo_voxel_keys": (_i, [_vp, _fl, _vp, _sz, _vp, _vp]),
    "lo_grid_create": (_i, [_vp, _fl, _sz, _sz, C.POINTER(_vp)]), "lo_grid_destroy": (None, [_vp]),
    "lo_grid_set_voxel_size": (_i, [_vp, _fl]), "lo_grid_set_max_points": (_i, [_vp, _sz]),
    "lo_grid_get_correspondence": (_i, [_vp, _vp, C.c_double, _vp, C.POINTER(_i)]),
    "lo_point_time_normalize": (_i, [_vp, _vp, _sz, _vp]), "lo_transform_non_rigid": (_i, [_vp, _vp, _sz, _vp, _vp, _vp]),
    "lo_classify_pair": (_i, [_vp, _vp, _sz, _vp, _sz, C.POINTER(_sz), _vp, _sz, C.POINTER(_sz)]), "lo_grid_size": (_i, [_vp, C.POINTER(_sz), C.POINTER(_sz)]),
    "lo_grid_add_cloud": (_i, [_vp, _vp, _sz, _i]), "lo_grid_add_cloud_dev": (_i, [_vp, _vp, _sz, _i]),
    "lo_grid_get_cloud": (_i, [_vp, _i, _i, _vp, _sz, C.POINTER(_sz)]),
    "lo_grid_radius_cleanup": (_i, [_vp, _vp, _fl]), "lo_robin_iteration_order": (_i, [_vp, _vp, _sz, _vp]),
    "lo_grid_snapshot": (_i, [_vp, _vp, _sz, C.POINTER(_sz)]), "lo_grid_restore": (_i, [_vp, _vp, _sz]),
    "lo_odometry_snapshot": (_i, [_vp, _vp, _sz, C.POINTER(_sz)]), "lo_odometry_restore": (_i, [_vp, _vp, _sz]),
    "lo_grid_find_matching_pairs": (_i, [_vp, _vp, _sz, _vp, _fl, _vp, _vp, C.POINTER(_sz)]),
    "lo_nn_residual": (_i, [_vp, _vp, _sz, _vp, _vp, _vp, _fl, _vp]),
    "lo_nn_residual_dev": (_i, [_vp, _vp, _sz, _vp, _vp, _vp, _fl, _vp]),
    "lo_nn_residual_static_dev": (_i, [_vp, _vp, _sz, _vp, _vp, _vp, _fl, _vp]),
    "lo_align": (_i, [_vp, _vp, _sz, _vp, _vp, _vp]), "lo_align_dev": (_i, [_vp, _vp, _sz, _vp, _vp, _vp]),
    "lo_grid_set_shard": (_i, [_vp, C.c_uint32, C.c_uint32]),
    "lo_shard_search_dev": (_i, [_vp, _vp, _sz, _vp, _fl, _vp, _vp]),
    "lo_shard_evaluate_dev": (_i, [_vp, _vp, _sz, _vp, _vp, _vp, _vp, _vp]),
    "lo_sharded_unique_id": (_i, [_vp]),
    "lo_sharded_create": (_i, [_vp, _vp, _i, _i, _fl, _sz, _sz, C.POINTER(_vp)]), "lo_sharded_destroy": (None, [_vp]),
    "lo_sharded_grid": (_vp, [_vp]), "lo_sharded_info_get": (_i, [_vp, _vp]),
    "lo_sharded_set_fused_sum": (_i, [_vp, _i]), "lo_sharded_set_fused_min": (_i, [_vp, _i]),
    "lo_sharded_create_group": (_i, [_vp, _i, _fl, _sz, _sz, _vp]),
    "lo_sharded_align_group": (_i, [_vp, _i, _vp, _sz, _vp, _fl, _vp, _vp]),
    "lo_sharded_align_group_host": (_i, [_vp, _i, _vp, _sz, _vp, _fl, _vp, _vp]),
    "lo_sharded_align": (_i, [_vp, _vp, _sz, _vp, _fl, _vp, _vp]), "lo_sharded_align_dev": (_i, [_vp, _vp, _sz, _vp, _fl, _vp, _vp]),
    "lo_lm_create": (_vp, []), "lo_lm_destroy": (None, [_vp]), "lo_lm_begin": (_i, [_vp, _vp, _vp, _vp]),
    "lo_lm_propose": (_i, [_vp, _vp]), "lo_lm_update": (_i, [_vp, _vp]), "lo_lm_result": (_i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "lo_deskew": (_i, [_vp, _vp, _sz, _vp, _vp, _vp]),
    "lo_transform_xyz": (_i, [_vp, _vp, _sz, _vp, _vp]), "lo_transform_with_normals": (_i, [_vp, _vp, _sz, _vp, _vp]),
    "lo_range_filter": (_i, [_vp, _vp, _sz, _fl, _fl, _vp, C.POINTER(_sz)]),
    "lo_classify": (_i, [_vp, _vp, _sz, _vp, _sz, C.POINTER(_sz)]),
    "lo_pca_normals": (_i, [_vp, _vp, _sz, _fl, _vp]),
    "lo_params_default": (None, [_vp]), "lo_params_from_yaml": (_i, [C.c_char_p, _vp]),
    "lo_odometry_create": (_i, [_vp, _vp, C.POINTER(_vp)]),
    "lo_odometry_create_sized": (_i, [_vp, _vp, _sz, _sz, C.POINTER(_vp)]), "lo_odometry_destroy": (None, [_vp]),
    "lo_odometry_process": (_i, [_vp, _vp, _sz]), "lo_odometry_process_dev": (_i, [_vp, _vp, _sz]),
    "lo_odometry_process_next": (_i, [_vp, _vp, _sz, _vp, _sz]),
    "lo_odometry_submit": (_i, [_vp, _vp, _sz]), "lo_odometry_submit_dev": (_i, [_vp, _vp, _sz]), "lo_odometry_wait": (_i, [_vp]),
    "lo_odometry_process_planar": (_i, [_vp, _vp, _sz]), "lo_odometry_pose": (_i, [_vp, _vp]),
    "lo_odometry_keyframe_info": (_i, [_vp, _vp]),
    "lo_odometry_host_threads": (_i, [_vp]),
    "lo_odometry_stats": (_i, [_vp, _vp]),
    "lo_odometry_keyframe": (_i, [_vp, _i, _vp, _sz, C.POINTER(_sz)]),
    "lo_odometry_keyframe_pn": (_i, [_vp, _vp, _sz, C.POINTER(_sz)]),
    "lo_odometry_deskewed": (_i, [_vp, _vp, _sz, C.POINTER(_sz)]),
}

_lib = None


def load_library():
    """Load liblo_b200.so (built in-tree by lidar_odometry_demo_b200/build.py). Fails loudly if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise LoError(LO_E_IO, f"{LIB_PATH} is missing: run `python lidar_odometry_demo_b200/build.py` "
                                   "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def _check(code):
    if code != LO_OK:
        raise LoError(code, load_library().lo_last_error().decode())


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _pose7(p):
    a = np.ascontiguousarray(p, dtype=np.float32)
    assert a.shape == (7,)
    return a


class Context:
    """One CUDA device + stream (lo_ctx)."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        _check(self.lib.lo_ctx_create(device, C.byref(h)))
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.lo_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    @property
    def stream(self):
        return self.lib.lo_ctx_stream(self.h)

    def synchronize(self):
        _check(self.lib.lo_ctx_synchronize(self.h))

    @property
    def kernel_launches(self):
        return int(self.lib.lo_ctx_kernel_launches(self.h))

    def robin_iteration_order(self, hash22):
        """Iteration order of the reference's tsl::robin_map after inserting distinct keys with these hashes in index order."""
        h = np.ascontiguousarray(hash22, dtype=np.uint32)
        out = np.empty(len(h), np.uint32)
        _check(self.lib.lo_robin_iteration_order(self.h, _p(h), len(h), _p(out)))
        return out

    def voxel_keys(self, voxel_size, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        keys = np.empty((len(cloud), 3), np.int64)
        h22 = np.empty(len(cloud), np.uint32)
        _check(self.lib.lo_voxel_keys(self.h, voxel_size, _p(cloud), len(cloud), _p(keys), _p(h22)))
        return keys, h22

    def deskew(self, cloud, start7, end7):
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        out = np.empty_like(cloud)
        _check(self.lib.lo_deskew(self.h, _p(cloud), len(cloud), _p(_pose7(start7)), _p(_pose7(end7)), _p(out)))
        return out

    def point_time_normalize(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        out = np.empty_like(cloud)
        _check(self.lib.lo_point_time_normalize(self.h, _p(cloud), len(cloud), _p(out)))
        return out

    def transform_non_rigid(self, cloud, start7, end7):
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        out = np.empty_like(cloud)
        _check(self.lib.lo_transform_non_rigid(self.h, _p(cloud), len(cloud), _p(_pose7(start7)), _p(_pose7(end7)), _p(out)))
        return out

    def classify_pair(self, cloud):
        """CloudClassifier::classify's pair: (planar with normals, unclassified)."""
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        planar, un = np.zeros(max(len(cloud), 1), PN), np.zeros(max(len(cloud), 1), XYZIRT)
        n, nu = C.c_size_t(0), C.c_size_t(0)
        _check(self.lib.lo_classify_pair(self.h, _p(cloud), len(cloud), _p(planar), len(planar), C.byref(n), _p(un), len(un), C.byref(nu)))
        return planar[:n.value].copy(), un[:nu.value].copy()

    def transform_xyz(self, cloud, pose7):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        out = np.empty_like(cloud)
        _check(self.lib.lo_transform_xyz(self.h, _p(cloud), len(cloud), _p(_pose7(pose7)), _p(out)))
        return out

    def transform_with_normals(self, cloud, pose7):
        cloud = np.ascontiguousarray(cloud, dtype=PN)
        out = np.empty_like(cloud)
        _check(self.lib.lo_transform_with_normals(self.h, _p(cloud), len(cloud), _p(_pose7(pose7)), _p(out)))
        return out

    def range_filter(self, cloud, mn, mx):
        cloud = np.ascontiguousarray(cloud, dtype=PN)
        out = np.empty_like(cloud)
        n = C.c_size_t(0)
        _check(self.lib.lo_range_filter(self.h, _p(cloud), len(cloud), mn, mx, _p(out), C.byref(n)))
        return out[:n.value].copy()

    def classify(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        out = np.zeros(max(len(cloud), 1), PN)
        n = C.c_size_t(0)
        _check(self.lib.lo_classify(self.h, _p(cloud), len(cloud), _p(out), len(out), C.byref(n)))
        return out[:n.value].copy()

    def pca_normals(self, cloud, radius):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        out = np.empty((len(cloud), 3), np.float32)
        _check(self.lib.lo_pca_normals(self.h, _p(cloud), len(cloud), radius, _p(out)))
        return out


class VoxelGrid:
    """Mirror of the reference's VoxelGrid (src/voxel_grid.h:17-257) over lo_grid."""

    def __init__(self, ctx, voxel_size=0.5, max_points=10, capacity_hint=0):
        self.ctx, self.lib = ctx, ctx.lib
        h = C.c_void_p()
        _check(self.lib.lo_grid_create(ctx.h, voxel_size, max_points, capacity_hint, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            self.lib.lo_grid_destroy(self.h)
        self.h = None

    def __del__(self):
        self.close()

    def setVoxelSize(self, voxel_size):
        _check(self.lib.lo_grid_set_voxel_size(self.h, voxel_size))

    def setMaxPoints(self, max_points):
        _check(self.lib.lo_grid_set_max_points(self.h, max_points))

    def getCorrespondence(self, query3, max_correspondence_distance_sq):
        """VoxelGrid::getCorrespondence: (valid, plane_origin[3], plane_normal[3]) as float64."""
        q = np.ascontiguousarray(query3, dtype=np.float32)
        pair, valid = Pair(), C.c_int(0)
        _check(self.lib.lo_grid_get_correspondence(self.h, _p(q), float(max_correspondence_distance_sq), C.byref(pair), C.byref(valid)))
        return bool(valid.value), np.array(pair.plane_origin), np.array(pair.plane_normal)

    def addCloud(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=PN)
        _check(self.lib.lo_grid_add_cloud(self.h, _p(cloud), len(cloud), 1))

    def addCloudWithoutNormals(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        _check(self.lib.lo_grid_add_cloud(self.h, _p(cloud), len(cloud), 0))

    def size(self):
        nv, npts = C.c_size_t(0), C.c_size_t(0)
        _check(self.lib.lo_grid_size(self.h, C.byref(nv), C.byref(npts)))
        return int(nv.value)

    def numPoints(self):
        nv, npts = C.c_size_t(0), C.c_size_t(0)
        _check(self.lib.lo_grid_size(self.h, C.byref(nv), C.byref(npts)))
        return int(npts.value)

    def _get(self, mode, order, dtype):
        n = C.c_size_t(0)
        _check(self.lib.lo_grid_get_cloud(self.h, mode, order, None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype)
        if n.value:
            _check(self.lib.lo_grid_get_cloud(self.h, mode, order, _p(out), len(out), C.byref(n)))
        return out

    def getCloud(self, reference_order=False):
        return self._get(0, int(reference_order), PN)

    def getCloudWithoutNormals(self, reference_order=False):
        return self._get(1, int(reference_order), XYZ)

    def getSparseCloudWithoutNormals(self, reference_order=False):
        return self._get(2, int(reference_order), XYZ)

    def snapshot(self):
        """The table as it lies in HBM (lo_grid_snapshot) as a uint8 array."""
        n = C.c_size_t(0)
        _check(self.lib.lo_grid_snapshot(self.h, None, 0, C.byref(n)))
        out = np.empty(n.value, np.uint8)
        _check(self.lib.lo_grid_snapshot(self.h, _p(out), len(out), C.byref(n)))
        return out

    def restore(self, blob):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        _check(self.lib.lo_grid_restore(self.h, _p(blob), len(blob)))

    def radiusCleanup(self, center, radius):
        c = np.ascontiguousarray(center, dtype=np.float32)
        _check(self.lib.lo_grid_radius_cleanup(self.h, _p(c), radius))

    def findMatchingPairs(self, cloud, pose7, max_dist=0.3):
        """Returns (pairs9 [m,9] float64: src, origin, normal; valid [m] bool), query order kept."""
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        m = len(cloud)
        pairs = np.zeros((m, 9), np.float64)
        valid = np.zeros(m, np.uint8)
        n = C.c_size_t(0)
        _check(self.lib.lo_grid_find_matching_pairs(self.h, _p(cloud), m, _p(_pose7(pose7)), max_dist, _p(pairs), _p(valid),
                                                    C.byref(n)))
        assert n.value == int(valid.sum())
        return pairs, valid.astype(bool)

    def nnResidual(self, cloud, pose7, x7d, prior_b, max_dist=0.3):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        x = np.ascontiguousarray(x7d, dtype=np.float64)
        b = np.ascontiguousarray(prior_b, dtype=np.float64)
        out = NormalEq()
        _check(self.lib.lo_nn_residual(self.h, _p(cloud), len(cloud), _p(_pose7(pose7)), _p(x), _p(b), max_dist, C.byref(out)))
        self.last_candidates, self.last_occupied = int(out.candidates), int(out.occupied)
        return out.cost, np.array(out.g[:]), np.array(out.jtj[:]).reshape(6, 6), int(out.n_valid)


class CloudMatcher:
    """Mirror of CloudMatcher (src/cloud_matcher.h:13-17)."""

    def align(self, keyframe, planar_cloud, position_guess7, with_stats=False):
        cloud = np.ascontiguousarray(planar_cloud, dtype=XYZ)
        out = np.empty(7, np.float32)
        stats = AlignStats()
        _check(keyframe.lib.lo_align(keyframe.h, _p(cloud), len(cloud), _p(_pose7(position_guess7)), _p(out), C.byref(stats)))
        return (out, stats) if with_stats else out


def sharded_unique_id():
    """ncclGetUniqueId through the library (rank 0 calls it and ships the 128 bytes to the other ranks)."""
    out = np.zeros(128, np.uint8)
    _check(load_library().lo_sharded_unique_id(_p(out)))
    return out


class ShardedCloudMatcher:
    """CloudMatcher::align over a key-frame map sharded by voxel-hash range, one process per GPU (lo_sharded_*): the
    collectives and the LM run inside the library, the host only enqueues."""

    def __init__(self, ctx, unique_id, rank, world, voxel_size=0.2, max_points=20, capacity_hint=0):
        self.ctx, self.lib = ctx, ctx.lib
        h = C.c_void_p()
        uid = np.ascontiguousarray(unique_id, np.uint8) if unique_id is not None else None
        _check(self.lib.lo_sharded_create(ctx.h, _p(uid), rank, world, voxel_size, max_points, capacity_hint, C.byref(h)))
        self.h = h
        self.grid = VoxelGrid.__new__(VoxelGrid)          # borrowed view of the rank's shard (owned by lo_sharded)
        self.grid.ctx, self.grid.lib, self.grid.h = ctx, ctx.lib, C.c_void_p(self.lib.lo_sharded_grid(h))
        self.grid.close = lambda: None

    def close(self):
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            self.grid.h = None
            self.lib.lo_sharded_destroy(self.h)
        self.h = None

    def __del__(self):
        self.close()

    def setFusedSum(self, enable):
        _check(self.lib.lo_sharded_set_fused_sum(self.h, int(bool(enable))))

    def setFusedMin(self, enable):
        _check(self.lib.lo_sharded_set_fused_min(self.h, int(bool(enable))))

    def info(self):
        out = ShardedInfo()
        _check(self.lib.lo_sharded_info_get(self.h, C.byref(out)))
        return out

    def align(self, planar_cloud, position_guess7, max_dist=0.3, with_stats=False):
        cloud = np.ascontiguousarray(planar_cloud, dtype=XYZ)
        out = np.empty(7, np.float32)
        stats = AlignStats()
        _check(self.lib.lo_sharded_align(self.h, _p(cloud), len(cloud), _p(_pose7(position_guess7)), max_dist, _p(out), C.byref(stats)))
        return (out, stats) if with_stats else out

    def alignDevice(self, device_ptr, m, position_guess7, max_dist=0.3):
        out = np.empty(7, np.float32)
        stats = AlignStats()
        _check(self.lib.lo_sharded_align_dev(self.h, C.c_void_p(device_ptr), m, _p(_pose7(position_guess7)), max_dist, _p(out), C.byref(stats)))
        return out, stats


class ShardedGroup:
    """The shards of one process (lo_sharded_create_group): one Context (stream) per shard, on one GPU or several; both
    exchanges run through peer memory, no NCCL. Driven together by align()."""

    def __init__(self, ctxs, voxel_size=0.2, max_points=20, capacity_hint=0):
        self.ctxs, self.lib, self.world = list(ctxs), ctxs[0].lib, len(ctxs)
        arr = (C.c_void_p * self.world)(*[c.h for c in self.ctxs])
        out = (C.c_void_p * self.world)()
        _check(self.lib.lo_sharded_create_group(arr, self.world, voxel_size, max_points, capacity_hint, out))
        self.h = [C.c_void_p(v) for v in out]
        self.grids = []
        for ctx, h in zip(self.ctxs, self.h):
            g = VoxelGrid.__new__(VoxelGrid)          # borrowed view of the shard (owned by lo_sharded)
            g.ctx, g.lib, g.h = ctx, ctx.lib, C.c_void_p(self.lib.lo_sharded_grid(h))
            g.close = lambda: None
            self.grids.append(g)

    def close(self):
        for g in getattr(self, "grids", []):
            g.h = None
        for ctx, h in zip(getattr(self, "ctxs", []), getattr(self, "h", [])):
            if h and getattr(ctx, "h", None):
                self.lib.lo_sharded_destroy(h)
        self.h = []

    def __del__(self):
        self.close()

    def addCloud(self, cloud):
        """The replicated key-frame cloud: every shard keeps the voxels it owns."""
        for g in self.grids:
            g.addCloud(cloud)

    def info(self, k=0):
        out = ShardedInfo()
        _check(self.lib.lo_sharded_info_get(self.h[k], C.byref(out)))
        return out

    def _handles(self):
        return (C.c_void_p * self.world)(*[h.value for h in self.h])

    def align(self, planar_cloud, position_guess7, max_dist=0.3):
        """Returns (poses [world, 7] — one per shard, all alike —, stats)."""
        cloud = np.ascontiguousarray(planar_cloud, dtype=XYZ)
        out = np.empty((self.world, 7), np.float32)
        stats = AlignStats()
        _check(self.lib.lo_sharded_align_group_host(self._handles(), self.world, _p(cloud), len(cloud), _p(_pose7(position_guess7)), max_dist,
                                                    _p(out), C.byref(stats)))
        return out, stats

    def alignDevice(self, device_ptrs, m, position_guess7, max_dist=0.3):
        ptrs = (C.c_void_p * self.world)(*device_ptrs)
        out = np.empty((self.world, 7), np.float32)
        stats = AlignStats()
        _check(self.lib.lo_sharded_align_group(self._handles(), self.world, ptrs, m, _p(_pose7(position_guess7)), max_dist, _p(out), C.byref(stats)))
        return out, stats


class LidarOdometry:
    """Mirror of LidarOdometry (src/lidar_odometry.h:18-86)."""

    def __init__(self, ctx, params=None, keyframe_voxels=0, scan_points=0):
        """keyframe_voxels / scan_points: capacities allocated once at creation (0: the library's defaults)."""
        self.ctx, self.lib = ctx, ctx.lib
        if params is None:
            params = Params()
            self.lib.lo_params_default(C.byref(params))
        self.params = params
        h = C.c_void_p()
        _check(self.lib.lo_odometry_create_sized(ctx.h, C.byref(params), keyframe_voxels, scan_points, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            self.lib.lo_odometry_destroy(self.h)
        self.h = None

    def __del__(self):
        self.close()

    def processCloud(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        _check(self.lib.lo_odometry_process(self.h, _p(cloud), len(cloud)))

    def processCloudHost(self, host_ptr, n):
        """processCloud on a host scan given by address (no numpy conversion): the drop-in call, lo_odometry_process."""
        _check(self.lib.lo_odometry_process(self.h, C.c_void_p(host_ptr), n))

    def processCloudWithNext(self, cloud_ptr, n, next_ptr, next_n):
        """processCloud on a host scan while the next one (pinned host memory) is uploaded behind it."""
        _check(self.lib.lo_odometry_process_next(self.h, C.c_void_p(cloud_ptr), n, C.c_void_p(next_ptr) if next_ptr else None, next_n))

    def submit(self, host_ptr, n):
        """Queue a host scan (lo_odometry_submit); at most QUEUE_DEPTH in flight."""
        _check(self.lib.lo_odometry_submit(self.h, C.c_void_p(host_ptr), n))

    def submitDevice(self, device_ptr, n):
        _check(self.lib.lo_odometry_submit_dev(self.h, C.c_void_p(device_ptr), n))

    def wait(self):
        """Block until the oldest queued scan is done; getCurrentPose() / stats() then describe it."""
        _check(self.lib.lo_odometry_wait(self.h))

    def processCloudDevice(self, device_ptr, n):
        _check(self.lib.lo_odometry_process_dev(self.h, C.c_void_p(device_ptr), n))

    def processPlanar(self, planar):
        planar = np.ascontiguousarray(planar, dtype=PN)
        _check(self.lib.lo_odometry_process_planar(self.h, _p(planar), len(planar)))

    def snapshot(self):
        """Checkpoint (poses + key-frame table) as a uint8 array: lo_odometry_snapshot."""
        n = C.c_size_t(0)
        _check(self.lib.lo_odometry_snapshot(self.h, None, 0, C.byref(n)))
        out = np.empty(n.value, np.uint8)
        _check(self.lib.lo_odometry_snapshot(self.h, _p(out), len(out), C.byref(n)))
        return out

    def restore(self, blob):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        _check(self.lib.lo_odometry_restore(self.h, _p(blob), len(blob)))

    def getCurrentPose(self):
        out = np.empty(7, np.float32)
        _check(self.lib.lo_odometry_pose(self.h, _p(out)))
        return out

    def stats(self):
        s = ScanStats()
        _check(self.lib.lo_odometry_stats(self.h, C.byref(s)))
        return s

    def hostThreads(self):
        """Host threads the blocking processCloud uses for a pageable scan (caller + staging-copy helpers)."""
        return int(self.lib.lo_odometry_host_threads(self.h))

    def keyframeInfo(self):
        k = KeyframeInfo()
        _check(self.lib.lo_odometry_keyframe_info(self.h, C.byref(k)))
        return k

    def _cloud(self, fn, dtype, *lead):
        n = C.c_size_t(0)
        _check(fn(self.h, *lead, None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype)
        if n.value:
            _check(fn(self.h, *lead, _p(out), len(out), C.byref(n)))
        return out

    def getKeyFrameCloud(self):
        return self._cloud(self.lib.lo_odometry_keyframe, XYZ, 0)

    def getFullKeyFrameCloud(self):
        return self._cloud(self.lib.lo_odometry_keyframe, XYZ, 1)

    def getKeyFrameCloudWithNormals(self):
        return self._cloud(self.lib.lo_odometry_keyframe_pn, PN)

    def getTempCloud(self):
        return self._cloud(self.lib.lo_odometry_deskewed, XYZIRT)


def params_from_yaml(path):
    p = Params()
    _check(load_library().lo_params_from_yaml(path.encode(), C.byref(p)))
    return p
