"""CPU stand-in for one rank's shard, built on the oracle port (test infrastructure): same interface as
lidar_odometry_demo_b200.sharded.GpuShard, so the host-side orchestration (ShardedCloudMatcher + collectives) can be
exercised with gloo on CPU."""
import numpy as np
import torch

from lidar_odometry_demo_b200.sharded import SIGN, shard_owner
from lidar_odometry_demo_b200 import synthetic as S


def hash22(keys3):
    k = keys3.astype(np.int64).view(np.uint64)
    h = (k[:, 0] * np.uint64(73856093)) ^ (k[:, 1] * np.uint64(19349669)) ^ (k[:, 2] * np.uint64(83492791))
    return (h & np.uint64((1 << 22) - 1)).astype(np.uint32)


class OracleShard:
    def __init__(self, port, voxel_size, max_points, rank, world):
        self.port, self.vs, self.rank, self.world = port, voxel_size, rank, world
        self.grid = port.grid(voxel_size, max_points)

    def add_cloud(self, pn):
        xyz = np.stack([pn["x"], pn["y"], pn["z"]], 1)
        keys = self.port.get_indices(self.vs, xyz)
        mine = shard_owner(keys, hash22(keys), self.world) == self.rank
        self.grid.add_cloud(pn[mine])

    def set_queries(self, xyz):
        self.q = np.ascontiguousarray(xyz, S.XYZ)

    def search(self, pose7, max_dist=0.3):
        pose7 = np.asarray(pose7, np.float32)
        self.pairs9, self.valid = self.grid.find_matching_pairs(self.q, pose7, max_dist)
        R = self.port.rotation_matrix(pose7).astype(np.float64)
        t = pose7[:3].astype(np.float64)
        p = np.stack([self.q["x"], self.q["y"], self.q["z"]], 1).astype(np.float64)
        wq = np.stack([((R[r, 0] * p[:, 0] + R[r, 1] * p[:, 1]) + R[r, 2] * p[:, 2]) + t[r] for r in range(3)], 1).astype(np.float32)
        o = self.pairs9[:, 3:6].astype(np.float32)
        d = wq - o
        d2 = d[:, 0] * d[:, 0] + (d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])
        kq, ko = self.port.get_indices(self.vs, wq), self.port.get_indices(self.vs, o)
        off = ko - kq + 1
        probe = (9 * off[:, 0] + 3 * off[:, 1] + off[:, 2]).astype(np.uint64)
        packed = (d2.view(np.uint32).astype(np.uint64) << np.uint64(32)) | (probe << np.uint64(24))
        packed[~self.valid] = np.uint64(0xFFFFFFFFFFFFFFFF)
        self.local = packed.view(np.int64) ^ SIGN
        return torch.from_numpy(self.local.copy())

    def evaluate(self, keys_global, x7d):
        mine = self.valid & (self.local == keys_global.numpy())
        x = np.asarray(x7d, np.float64)
        acc = np.zeros(32)
        # the oracle's evaluate always adds the NormalPrior: neutralise it (b = t makes its residual 0; its 100*I is removed)
        cost, g, jtj = self.port.evaluate(self.pairs9[mine], x, x[4:7])
        jtj = jtj.copy()
        for k in range(3, 6):
            jtj[k, k] -= 100.0
        at = 0
        for a in range(6):
            for b in range(a, 6):
                acc[at] = jtj[a, b]
                at += 1
        acc[21:27], acc[27], acc[28] = g, cost, float(mine.sum())
        return torch.from_numpy(acc)
