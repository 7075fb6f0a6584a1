"""Helpers shared by the test modules."""
import numpy as np

from lidar_odometry_demo_b200 import synthetic as S

PN, XYZ, XYZIRT = S.PN, S.XYZ, S.XYZIRT


def bits_equal(a, b, names):
    return all(np.array_equal(np.ascontiguousarray(a[k]).view(np.uint32), np.ascontiguousarray(b[k]).view(np.uint32))
               for k in names)


def random_pn(n, seed, spread=20.0):
    rng = np.random.default_rng(seed)
    out = np.zeros(n, PN)
    p = (rng.normal(size=(n, 3)) * spread).astype(np.float32)
    nv = rng.normal(size=(n, 3))
    nv /= np.linalg.norm(nv, axis=1, keepdims=True)
    out["x"], out["y"], out["z"] = p.T
    out["nx"], out["ny"], out["nz"] = nv.astype(np.float32).T
    out["pad0"] = 1.0
    return out


def pn_to_xyz(pn):
    out = np.zeros(len(pn), XYZ)
    out["x"], out["y"], out["z"], out["pad"] = pn["x"], pn["y"], pn["z"], 1.0
    return out


def sort_rows(pn, names=("x", "y", "z", "nx", "ny", "nz")):
    """Canonical order for comparing clouds as multisets."""
    key = np.stack([np.ascontiguousarray(pn[k]).view(np.uint32) for k in names if k in pn.dtype.names], 1)
    order = np.lexsort(key.T[::-1])
    return key[order]


def planar_scene(port, scans, k):
    """Planar cloud of scan k as the oracle's classifier sees it (deskew with identity motion)."""
    ident = np.array([0, 0, 0, 1, 0, 0, 0], np.float32)
    d = port.transform_nonrigid(port.time_normalize(scans[k]), ident, ident)
    planar, _ = port.classify(d)
    return port.range_filter(planar, 4.0, 80.0)


def pose_error(a, b):
    """translation distance [m], rotation angle [rad] between two pose7."""
    dt = float(np.linalg.norm(np.asarray(a[:3], np.float64) - np.asarray(b[:3], np.float64)))
    qa, qb = np.asarray(a[3:], np.float64), np.asarray(b[3:], np.float64)
    qa, qb = qa / np.linalg.norm(qa), qb / np.linalg.norm(qb)
    d = min(1.0, abs(float(qa @ qb)))
    return dt, 2.0 * np.arccos(d)
