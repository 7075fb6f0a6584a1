"""Synthetic LiDAR workloads of BASELINE.json configs 2-5: an analytic street-canyon
scene ray-cast from a moving spinning LiDAR, written out as lidar_point::PointXYZIRT records
(/root/reference/src/lidar_point_type.h:13-31, 32 B). Host-side numpy only: workload generation is not
part of the timed path.
"""
import numpy as np

XYZIRT = np.dtype([("x", "f4"), ("y", "f4"), ("z", "f4"), ("pad0", "f4"), ("intensity", "f4"), ("ring", "u2"),
                   ("pad1", "u2"), ("time", "f4"), ("pad2", "f4")])
XYZ = np.dtype([("x", "f4"), ("y", "f4"), ("z", "f4"), ("pad", "f4")])
PN = np.dtype([("x", "f4"), ("y", "f4"), ("z", "f4"), ("pad0", "f4"), ("nx", "f4"), ("ny", "f4"), ("nz", "f4"),
               ("pad1", "f4"), ("curvature", "f4"), ("pad2", "f4", (3,))])


class Scene:
    """Ground plane z=-1.8, canyon walls y=+-half_width, axis-aligned boxes and vertical cylinders."""

    def __init__(self, seed=1234, length=600.0, half_width=12.0, n_boxes=None, n_cyl=None, scale=1.0):
        rng = np.random.default_rng(seed)
        self.ground_z = -1.8
        self.half_width = half_width * scale
        n_boxes = int(length / 2.5) if n_boxes is None else n_boxes
        n_cyl = int(length / 5) if n_cyl is None else n_cyl
        cx = rng.uniform(-20, length, n_boxes)
        side = rng.choice([-1.0, 1.0], n_boxes)
        cy = side * rng.uniform(4.0, self.half_width - 1.5, n_boxes)
        sx, sy, sz = rng.uniform(1.0, 4.0, n_boxes), rng.uniform(1.0, 3.0, n_boxes), rng.uniform(1.0, 6.0, n_boxes)
        self.box_lo = np.stack([cx - sx, cy - sy, np.full(n_boxes, self.ground_z)], 1)
        self.box_hi = np.stack([cx + sx, cy + sy, self.ground_z + sz], 1)
        self.cyl_c = np.stack([rng.uniform(-20, length, n_cyl),
                               rng.choice([-1.0, 1.0], n_cyl) * rng.uniform(3.0, self.half_width - 1.0, n_cyl)], 1)
        self.cyl_r = rng.uniform(0.15, 0.6, n_cyl)
        self.cyl_top = self.ground_z + rng.uniform(2.0, 8.0, n_cyl)

    def cast(self, o, d, max_range=120.0, cull=90.0):
        """o, d: (n,3) float64 ray origins / unit directions -> range (n,), inf where nothing is hit."""
        n = len(o)
        t = np.full(n, np.inf)
        with np.errstate(divide="ignore", invalid="ignore"):
            tg = (self.ground_z - o[:, 2]) / d[:, 2]
            t = np.where((tg > 0) & (tg < t), tg, t)
            for s in (-1.0, 1.0):
                tw = (s * self.half_width - o[:, 1]) / d[:, 1]
                hz = o[:, 2] + tw * d[:, 2]
                ok = (tw > 0) & (tw < t) & (hz < self.ground_z + 15.0)
                t = np.where(ok, tw, t)
            # boxes near the sensor only (slab test)
            cx = o[:, 0].mean()
            sel = np.nonzero((self.box_hi[:, 0] > cx - cull) & (self.box_lo[:, 0] < cx + cull))[0]
            inv = 1.0 / d
            for b in sel:
                t0 = (self.box_lo[b] - o) * inv
                t1 = (self.box_hi[b] - o) * inv
                tn = np.minimum(t0, t1).max(1)
                tf = np.maximum(t0, t1).min(1)
                ok = (tn <= tf) & (tn > 0) & (tn < t)
                t = np.where(ok, tn, t)
            sel = np.nonzero(np.abs(self.cyl_c[:, 0] - cx) < cull)[0]
            for c in sel:
                ox, oy = o[:, 0] - self.cyl_c[c, 0], o[:, 1] - self.cyl_c[c, 1]
                a = d[:, 0] ** 2 + d[:, 1] ** 2
                bq = ox * d[:, 0] + oy * d[:, 1]
                cq = ox ** 2 + oy ** 2 - self.cyl_r[c] ** 2
                disc = bq * bq - a * cq
                tc = (-bq - np.sqrt(np.maximum(disc, 0))) / a
                hz = o[:, 2] + tc * d[:, 2]
                ok = (disc > 0) & (tc > 0) & (tc < t) & (hz < self.cyl_top[c]) & (hz > self.ground_z)
                t = np.where(ok, tc, t)
        return t


def _quat_mul(a, b):
    aw, ax, ay, az = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bw, bx, by, bz = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx], -1)


def _rot(q, v):
    w, x, y, z = q[..., 0:1], q[..., 1:2], q[..., 2:3], q[..., 3:4]
    qv = np.concatenate([x, y, z], -1)
    uv = 2.0 * np.cross(qv, v)
    return v + w * uv + np.cross(qv, uv)


def trajectory(n_poses, seed=99, dt=0.1):
    """Positions and yaw for n_poses at 10 Hz: speed 2-8 m/s along +x, |yaw rate| <= 20 deg/s."""
    rng = np.random.default_rng(seed)
    t = np.arange(n_poses) * dt
    speed = (5.0 + 3.0 * np.sin(0.05 * t + rng.uniform(0, 6.28))) * np.minimum(1.0, t / 3.0)  # starts from rest
    yaw = np.deg2rad(6.0) * np.sin(0.35 * t + rng.uniform(0, 6.28))  # peak rate ~2.1 deg/s * ... < 20 deg/s
    pos = np.zeros((n_poses, 3))
    for k in range(1, n_poses):
        pos[k, 0] = pos[k - 1, 0] + speed[k - 1] * np.cos(yaw[k - 1]) * dt
        pos[k, 1] = pos[k - 1, 1] + speed[k - 1] * np.sin(yaw[k - 1]) * dt
    pos[:, 1] = np.clip(pos[:, 1], -2.5, 2.5)
    q = np.stack([np.cos(yaw / 2), np.zeros_like(yaw), np.zeros_like(yaw), np.sin(yaw / 2)], -1)
    return pos, q, yaw


class LidarModel:
    def __init__(self, rings=16, az_steps=1875, elev_min=-15.0, elev_max=15.0):
        self.rings, self.az_steps = rings, az_steps
        self.elev = np.deg2rad(np.linspace(elev_min, elev_max, rings))

    @staticmethod
    def vlp16():
        return LidarModel(16, 1875, -15.0, 15.0)

    @staticmethod
    def beam128():
        return LidarModel(128, 2048, -25.0, 15.0)


def generate_scan(scene, lidar, pos0, q0, pos1, q1, seed, noise=0.02, t_offset=0.0, sweep=0.1):
    """One sweep while the sensor moves from (pos0,q0) to (pos1,q1). Points are in the sensor frame at
    their own firing time (i.e. skewed), ordered azimuth-major like a spinning sensor's packets."""
    rng = np.random.default_rng(seed)
    A, R = lidar.az_steps, lidar.rings
    az_idx = np.repeat(np.arange(A), R)
    ring = np.tile(np.arange(R), A)
    s = az_idx / A
    az = 2 * np.pi * s
    el = lidar.elev[ring]
    d_s = np.stack([np.cos(el) * np.cos(az), -np.cos(el) * np.sin(az), np.sin(el)], 1)  # clockwise spin
    yaw0 = 2 * np.arctan2(q0[3], q0[0])
    yaw1 = 2 * np.arctan2(q1[3], q1[0])
    yaw = yaw0 + (yaw1 - yaw0) * s
    q = np.stack([np.cos(yaw / 2), np.zeros_like(yaw), np.zeros_like(yaw), np.sin(yaw / 2)], -1)
    o = pos0[None, :] + (pos1 - pos0)[None, :] * s[:, None]
    d_w = _rot(q, d_s)
    rng_t = scene.cast(o, d_w)
    rng_t = rng_t + rng.normal(0.0, noise, len(rng_t))
    keep = np.isfinite(rng_t) & (rng_t > 0.5) & (rng_t < 120.0)
    out = np.zeros(int(keep.sum()), XYZIRT)
    p = d_s[keep] * rng_t[keep, None]
    out["x"], out["y"], out["z"] = p[:, 0], p[:, 1], p[:, 2]
    out["intensity"] = rng.uniform(0, 100, len(out))
    out["ring"] = ring[keep]
    out["time"] = (t_offset + s[keep] * sweep).astype(np.float32)
    return out


_POOL_JOB = None   # (scene, lidar, pos, q, scene_seed) of the stream being generated, inherited by forked workers


def _pool_scan(k):
    scene, lidar, pos, q, scene_seed = _POOL_JOB
    return generate_scan(scene, lidar, pos[k], q[k], pos[k + 1], q[k + 1], scene_seed + k, t_offset=0.1 * k)


def vlp16_stream(n_scans, scene_seed=1234, traj_seed=99, lidar=None, scene=None, workers=1):
    """Config 2: list of XYZIRT scans plus the ground-truth poses at sweep start. The scene's length (and with it the
    boxes drawn from the seed) follows n_scans, so streams of different lengths are different streams. workers > 1:
    the scans (independent given the trajectory) are generated by forked worker processes — call it before CUDA is
    initialised in this process."""
    global _POOL_JOB
    lidar = lidar or LidarModel.vlp16()
    pos, q, _ = trajectory(n_scans + 1, traj_seed)
    scene = scene or Scene(scene_seed, length=max(60.0, pos[-1, 0] + 140.0))
    if workers > 1 and n_scans > 1:
        import multiprocessing as mp
        _POOL_JOB = (scene, lidar, pos, q, scene_seed)
        try:
            with mp.get_context("fork").Pool(min(workers, n_scans)) as pool:
                scans = pool.map(_pool_scan, range(n_scans), chunksize=max(1, n_scans // (4 * workers)))
        finally:
            _POOL_JOB = None
        return scans, pos, q
    scans = [generate_scan(scene, lidar, pos[k], q[k], pos[k + 1], q[k + 1], scene_seed + k, t_offset=0.1 * k)
             for k in range(n_scans)]
    return scans, pos, q


def planar_patch_map(n_points, seed=777, extent=(1000.0, 1000.0, 40.0), patch=2.0, pts_per_patch=2000):
    """Config 4: key-frame points with unit normals on random planar patches (PointNormal records)."""
    rng = np.random.default_rng(seed)
    n_patches = max(1, n_points // pts_per_patch)
    out = np.zeros(n_patches * pts_per_patch, PN)
    c = rng.uniform([0, 0, 0], extent, (n_patches, 3))
    nrm = rng.normal(size=(n_patches, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    a = np.cross(nrm, rng.normal(size=(n_patches, 3)))
    a /= np.linalg.norm(a, axis=1, keepdims=True)
    b = np.cross(nrm, a)
    uv = rng.uniform(-patch / 2, patch / 2, (n_patches, pts_per_patch, 2))
    p = c[:, None, :] + uv[..., 0:1] * a[:, None, :] + uv[..., 1:2] * b[:, None, :]
    p = p.reshape(-1, 3).astype(np.float32)
    out["x"], out["y"], out["z"] = p[:, 0], p[:, 1], p[:, 2]
    n3 = np.repeat(nrm, pts_per_patch, 0).astype(np.float32)
    out["nx"], out["ny"], out["nz"] = n3[:, 0], n3[:, 1], n3[:, 2]
    out["pad0"] = 1.0
    return out


def to_xyz(points3):
    points3 = np.asarray(points3, dtype=np.float32).reshape(-1, 3)
    out = np.zeros(len(points3), XYZ)
    out["x"], out["y"], out["z"], out["pad"] = points3[:, 0], points3[:, 1], points3[:, 2], 1.0
    return out


def read_pcd_xyz(path):
    """Binary PCD v0.7 reader for the reference's fixture layout (FIELDS rgb _ x y z _, 32 B stride)."""
    with open(path, "rb") as f:
        fields = sizes = counts = None
        npts = 0
        while True:
            line = f.readline().decode("ascii", "replace").strip()
            tok = line.split()
            if not tok:
                continue
            if tok[0] == "FIELDS":
                fields = tok[1:]
            elif tok[0] == "SIZE":
                sizes = [int(v) for v in tok[1:]]
            elif tok[0] == "COUNT":
                counts = [int(v) for v in tok[1:]]
            elif tok[0] == "POINTS":
                npts = int(tok[1])
            elif tok[0] == "DATA":
                if tok[1] != "binary":
                    raise ValueError("only binary PCD supported")
                break
        offs, o = {}, 0
        for name, s, c in zip(fields, sizes, counts):
            offs.setdefault(name, o)
            o += s * c
        raw = np.frombuffer(f.read(o * npts), dtype=np.uint8).reshape(npts, o)
    cols = [raw[:, offs[k]:offs[k] + 4].copy().view(np.float32)[:, 0] for k in ("x", "y", "z")]
    return np.stack(cols, 1)
