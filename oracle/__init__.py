"""TEST INFRASTRUCTURE: ctypes bindings for the two CPU checkers (see oracle/oracle_api.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package (lidar_odometry_demo_b200) never does.

    load("port")       -> oracle/liblo_oracle.so   (restatement, builds anywhere: make -C oracle port)
    load("reference")  -> oracle/_ref/liblo_ref.so (reference's own headers; built where /root/reference exists)
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
XYZ = np.dtype([("x", "f4"), ("y", "f4"), ("z", "f4"), ("pad", "f4")])
PN = np.dtype([("x", "f4"), ("y", "f4"), ("z", "f4"), ("pad0", "f4"), ("nx", "f4"), ("ny", "f4"), ("nz", "f4"),
               ("pad1", "f4"), ("curvature", "f4"), ("pad2", "f4", (3,))])
XYZIRT = np.dtype([("x", "f4"), ("y", "f4"), ("z", "f4"), ("pad0", "f4"), ("intensity", "f4"), ("ring", "u2"),
                   ("pad1", "u2"), ("time", "f4"), ("pad2", "f4")])
assert XYZ.itemsize == 16 and PN.itemsize == 48 and XYZIRT.itemsize == 32
ALIGN_LOG_STRIDE = 16


class Params(C.Structure):
    _fields_ = [("lidar_min_range", C.c_float), ("lidar_max_range", C.c_float), ("keyframe_voxel_size", C.c_float),
                ("keyframe_max_points_cnt", C.c_uint64), ("keyframe_matching_voxel_size", C.c_float),
                ("keyframe_update_voxel_size", C.c_float), ("keyframe_cleanup_range", C.c_float),
                ("angular_divergence_threshold", C.c_float)]


def build(kind="port"):
    target = "port" if kind == "port" else "ref"
    subprocess.run(["make", "-C", _HERE, target], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)


def lib_path(kind):
    return os.path.join(_HERE, "liblo_oracle.so") if kind == "port" else os.path.join(_HERE, "_ref", "liblo_ref.so")


def available(kind):
    return os.path.exists(lib_path(kind))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f7(p):
    a = np.ascontiguousarray(p, dtype=np.float32)
    assert a.shape == (7,)
    return a


class Oracle:
    def __init__(self, kind="port"):
        path = lib_path(kind)
        if not os.path.exists(path):
            build(kind)
        L = self.L = C.CDLL(path)
        vp, sz, fl = C.c_void_p, C.c_size_t, C.c_float
        sig = {
            "orc_kind": (C.c_char_p, []),
            "orc_grid_create": (vp, [fl, sz]), "orc_grid_clone": (vp, [vp]), "orc_grid_destroy": (None, [vp]),
            "orc_grid_add_cloud": (None, [vp, vp, sz]), "orc_grid_add_cloud_xyz": (None, [vp, vp, sz]),
            "orc_grid_size": (sz, [vp]), "orc_grid_num_points": (sz, [vp]),
            "orc_grid_get_cloud": (sz, [vp, vp, sz]), "orc_grid_get_cloud_xyz": (sz, [vp, vp, sz]),
            "orc_grid_get_sparse_xyz": (sz, [vp, vp, sz]), "orc_grid_radius_cleanup": (None, [vp, vp, fl]),
            "orc_grid_set_max_points": (None, [vp, sz]), "orc_grid_get_correspondence": (C.c_int, [vp, vp, C.c_double, vp, vp]),
            "orc_get_indices": (None, [fl, vp, sz, sz, vp]),
            "orc_hash": (C.c_uint64, [C.c_int64, C.c_int64, C.c_int64]),
            "orc_grid_find_matching_pairs": (sz, [vp, vp, sz, vp, fl, vp, vp]),
            "orc_time_normalize": (None, [vp, sz, vp]), "orc_transform_nonrigid": (None, [vp, sz, vp, vp, vp]),
            "orc_transform_xyz": (None, [vp, sz, vp, vp]), "orc_transform_with_normals": (None, [vp, sz, vp, vp]),
            "orc_range_filter": (sz, [vp, sz, fl, fl, vp]), "orc_classify": (sz, [vp, sz, vp, vp]), "orc_classify_full": (sz, [vp, sz, vp, vp, vp]),
            "orc_pose_compose": (None, [vp, vp, vp]), "orc_pose_inverse": (None, [vp, vp]),
            "orc_pose_relative_to": (None, [vp, vp, vp]), "orc_pose_rotation_matrix": (None, [vp, vp]),
            "orc_evaluate": (None, [vp, sz, vp, vp, vp, vp, vp, vp, vp]),
            "orc_solve": (C.c_int, [vp, sz, vp, vp, vp, C.c_int, vp]),
            "orc_align": (C.c_int, [vp, vp, sz, vp, vp, vp]),
            "orc_pca_normals": (None, [vp, sz, fl, vp]),
            "orc_params_default": (None, [vp]), "orc_odom_create": (vp, [vp]), "orc_odom_destroy": (None, [vp]),
            "orc_odom_process": (None, [vp, vp, sz, vp, vp]), "orc_odom_process_planar": (None, [vp, vp, sz, vp, vp]), "orc_odom_pose": (None, [vp, vp]),
            "orc_odom_keyframe": (sz, [vp, C.c_int, vp, sz]), "orc_odom_keyframe_pn": (sz, [vp, vp, sz]),
            "orc_odom_deskewed": (sz, [vp, vp, sz]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        self.kind = L.orc_kind().decode()

    # ---- grid ----
    def grid(self, voxel_size, max_points):
        return OracleGrid(self, self.L.orc_grid_create(voxel_size, max_points))

    def get_indices(self, voxel_size, xyz):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
        out = np.empty((len(xyz), 3), np.int64)
        self.L.orc_get_indices(voxel_size, _p(xyz), 3, len(xyz), _p(out))
        return out

    def hash(self, ix, iy, iz):
        return int(self.L.orc_hash(int(ix), int(iy), int(iz)))

    # ---- utilities ----
    def time_normalize(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        out = np.empty_like(cloud)
        self.L.orc_time_normalize(_p(cloud), len(cloud), _p(out))
        return out

    def transform_nonrigid(self, cloud, start7, end7):
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        out = np.empty_like(cloud)
        self.L.orc_transform_nonrigid(_p(cloud), len(cloud), _p(_f7(start7)), _p(_f7(end7)), _p(out))
        return out

    def transform_xyz(self, cloud, pose7):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        out = np.empty_like(cloud)
        self.L.orc_transform_xyz(_p(cloud), len(cloud), _p(_f7(pose7)), _p(out))
        return out

    def transform_with_normals(self, cloud, pose7):
        cloud = np.ascontiguousarray(cloud, dtype=PN)
        out = np.empty_like(cloud)
        self.L.orc_transform_with_normals(_p(cloud), len(cloud), _p(_f7(pose7)), _p(out))
        return out

    def range_filter(self, cloud, mn, mx):
        cloud = np.ascontiguousarray(cloud, dtype=PN)
        out = np.empty_like(cloud)
        n = self.L.orc_range_filter(_p(cloud), len(cloud), mn, mx, _p(out))
        return out[:n].copy()

    def classify(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        out = np.zeros(max(len(cloud), 1), PN)
        nun = C.c_size_t(0)
        n = self.L.orc_classify(_p(cloud), len(cloud), _p(out), C.byref(nun))
        return out[:n].copy(), int(nun.value)

    def classify_full(self, cloud):
        """CloudClassifier::classify's pair: (planar with normals, unclassified points with intensity = curvature)."""
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        out, un = np.zeros(max(len(cloud), 1), PN), np.zeros(max(len(cloud), 1), XYZIRT)
        nun = C.c_size_t(0)
        n = self.L.orc_classify_full(_p(cloud), len(cloud), _p(out), _p(un), C.byref(nun))
        return out[:n].copy(), un[:nun.value].copy()

    def pose_compose(self, a, b):
        out = np.empty(7, np.float32)
        self.L.orc_pose_compose(_p(_f7(a)), _p(_f7(b)), _p(out))
        return out

    def pose_inverse(self, a):
        out = np.empty(7, np.float32)
        self.L.orc_pose_inverse(_p(_f7(a)), _p(out))
        return out

    def pose_relative_to(self, a, b):
        out = np.empty(7, np.float32)
        self.L.orc_pose_relative_to(_p(_f7(a)), _p(_f7(b)), _p(out))
        return out

    def rotation_matrix(self, a):
        out = np.empty(9, np.float32)
        self.L.orc_pose_rotation_matrix(_p(_f7(a)), _p(out))
        return out.reshape(3, 3)

    # ---- solver ----
    def evaluate(self, pairs9, x7d, prior_b, want_rows=False):
        pairs9 = np.ascontiguousarray(pairs9, dtype=np.float64).reshape(-1, 9)
        n = len(pairs9)
        x = np.ascontiguousarray(x7d, dtype=np.float64)
        b = np.ascontiguousarray(prior_b, dtype=np.float64)
        cost = C.c_double(0)
        g, jtj = np.empty(6), np.empty(36)
        r = np.empty(n) if want_rows else None
        J = np.empty((n, 6)) if want_rows else None
        self.L.orc_evaluate(_p(pairs9), n, _p(x), _p(b), C.byref(cost), _p(g), _p(jtj),
                            _p(r) if want_rows else None, _p(J) if want_rows else None)
        return (cost.value, g, jtj.reshape(6, 6)) + ((r, J) if want_rows else ())

    def solve(self, pairs9, x7d, prior_b):
        pairs9 = np.ascontiguousarray(pairs9, dtype=np.float64).reshape(-1, 9)
        x = np.array(x7d, dtype=np.float64)
        b = np.ascontiguousarray(prior_b, dtype=np.float64)
        log = np.zeros((16, 8))
        term = C.c_int(0)
        k = self.L.orc_solve(_p(pairs9), len(pairs9), _p(b), _p(x), _p(log), 16, C.byref(term))
        return x, log[:k], term.value

    def pca_normals(self, cloud, radius):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        out = np.empty((len(cloud), 3), np.float32)
        self.L.orc_pca_normals(_p(cloud), len(cloud), radius, _p(out))
        return out

    def odometry(self, params=None):
        return OracleOdometry(self, params)

    def default_params(self):
        p = Params()
        self.L.orc_params_default(C.byref(p))
        return p


class OracleGrid:
    def __init__(self, orc, handle):
        self.o, self.h = orc, handle

    def __del__(self):
        if getattr(self, "h", None):
            self.o.L.orc_grid_destroy(self.h)
            self.h = None

    def clone(self):
        return OracleGrid(self.o, self.o.L.orc_grid_clone(self.h))

    def add_cloud(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=PN)
        self.o.L.orc_grid_add_cloud(self.h, _p(cloud), len(cloud))

    def add_cloud_xyz(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        self.o.L.orc_grid_add_cloud_xyz(self.h, _p(cloud), len(cloud))

    def size(self):
        return int(self.o.L.orc_grid_size(self.h))

    def num_points(self):
        return int(self.o.L.orc_grid_num_points(self.h))

    def get_cloud(self):
        out = np.zeros(self.num_points(), PN)
        n = self.o.L.orc_grid_get_cloud(self.h, _p(out), len(out))
        assert n == len(out)
        return out

    def get_cloud_xyz(self):
        out = np.zeros(self.num_points(), XYZ)
        n = self.o.L.orc_grid_get_cloud_xyz(self.h, _p(out), len(out))
        assert n == len(out)
        return out

    def get_sparse_xyz(self):
        out = np.zeros(self.size(), XYZ)
        n = self.o.L.orc_grid_get_sparse_xyz(self.h, _p(out), len(out))
        assert n == len(out)
        return out

    def radius_cleanup(self, center, radius):
        c = np.ascontiguousarray(center, dtype=np.float32)
        self.o.L.orc_grid_radius_cleanup(self.h, _p(c), radius)

    def set_max_points(self, max_points):
        self.o.L.orc_grid_set_max_points(self.h, max_points)

    def get_correspondence(self, query3, max_sq):
        q = np.ascontiguousarray(query3, dtype=np.float32)
        o, nv = np.zeros(3), np.zeros(3)
        ok = self.o.L.orc_grid_get_correspondence(self.h, _p(q), float(max_sq), _p(o), _p(nv))
        return bool(ok), o, nv

    def find_matching_pairs(self, cloud, pose7, max_dist=0.3):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        out9 = np.zeros((len(cloud), 9))
        valid = np.zeros(len(cloud), np.uint8)
        self.o.L.orc_grid_find_matching_pairs(self.h, _p(cloud), len(cloud), _p(_f7(pose7)), max_dist, _p(out9), _p(valid))
        return out9, valid.astype(bool)

    def align(self, cloud, guess7):
        cloud = np.ascontiguousarray(cloud, dtype=XYZ)
        out = np.empty(7, np.float32)
        log = np.zeros((35, ALIGN_LOG_STRIDE))
        k = self.o.L.orc_align(self.h, _p(cloud), len(cloud), _p(_f7(guess7)), _p(out), _p(log))
        return out, log[:k]


class OracleOdometry:
    def __init__(self, orc, params=None):
        self.o = orc
        self.params = params if params is not None else orc.default_params()
        self.h = orc.L.orc_odom_create(C.byref(self.params))

    def __del__(self):
        if getattr(self, "h", None):
            self.o.L.orc_odom_destroy(self.h)
            self.h = None

    def process(self, cloud):
        cloud = np.ascontiguousarray(cloud, dtype=XYZIRT)
        counts = np.zeros(4, np.uint64)
        stats = np.zeros(4)
        self.o.L.orc_odom_process(self.h, _p(cloud), len(cloud), _p(counts), _p(stats))
        return counts, stats

    def process_planar(self, planar):
        planar = np.ascontiguousarray(planar, dtype=PN)
        counts = np.zeros(4, np.uint64)
        stats = np.zeros(4)
        self.o.L.orc_odom_process_planar(self.h, _p(planar), len(planar), _p(counts), _p(stats))
        return counts, stats

    def pose(self):
        out = np.empty(7, np.float32)
        self.o.L.orc_odom_pose(self.h, _p(out))
        return out

    def keyframe_pn(self):
        n = self.o.L.orc_odom_keyframe_pn(self.h, None, 0)
        out = np.zeros(n, PN)
        self.o.L.orc_odom_keyframe_pn(self.h, _p(out), n)
        return out

    def keyframe(self, full=False):
        n = self.o.L.orc_odom_keyframe(self.h, int(full), None, 0)
        out = np.zeros(n, XYZ)
        self.o.L.orc_odom_keyframe(self.h, int(full), _p(out), n)
        return out

    def deskewed(self):
        n = self.o.L.orc_odom_deskewed(self.h, None, 0)
        out = np.zeros(n, XYZIRT)
        self.o.L.orc_odom_deskewed(self.h, _p(out), n)
        return out


_cache = {}


def load(kind="port"):
    if kind not in _cache:
        _cache[kind] = Oracle(kind)
    return _cache[kind]
