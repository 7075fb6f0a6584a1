"""CloudMatcher::align against a key-frame map sharded by voxel-hash range (BASELINE.json config 4).

Host-side orchestration of the per-rank phases of include/lo_b200.h (lo_shard_search_dev, lo_shard_evaluate_dev,
lo_lm_*): per evaluation one all-reduce(MIN) over the packed (d2, probe, slot) keys of the M replicated queries and
one all-reduce(SUM) over 32 doubles (21 JtJ + 6 Jtr + cost + count). The loop mirrors
/root/reference/src/cloud_matcher.cpp:105-178 (35 outer iterations, float round trip of the pose, stop rule on the
last recorded step norm). A C++ caller replaces the two collectives below by ncclAllReduce.

A process may hold several shards (a list of backends): they are reduced locally first, then across processes with
torch.distributed (NCCL on GPUs; gloo in the CPU test, where the backend is built on the oracle instead of CUDA).
"""
import ctypes as C

import numpy as np

SIGN = np.int64(-2 ** 63)


def shard_owner(keys3, hash22, world):
    """Rank that owns each voxel: the reference's 22-bit hash cut into `world` equal ranges (csrc/lo_device.cuh)."""
    return ((hash22.astype(np.uint64) * np.uint64(world)) >> np.uint64(22)).astype(np.int64)


class GpuShard:
    """One rank's shard on a GPU: lo_grid with lo_grid_set_shard + the replicated queries resident in HBM."""

    def __init__(self, lo, ctx, voxel_size, max_points, rank, world, capacity_hint=0):
        import torch
        self.lo, self.ctx, self.torch = lo, ctx, torch
        self.grid = lo.VoxelGrid(ctx, voxel_size, max_points, capacity_hint)
        lo._check(ctx.lib.lo_grid_set_shard(self.grid.h, rank, world))
        self.dev = torch.device("cuda", ctx.device)
        self.stream = torch.cuda.ExternalStream(ctx.stream, device=self.dev)

    def add_cloud(self, pn):            # replicated update cloud: only the owned voxels are kept
        self.grid.addCloud(pn)

    def set_queries(self, xyz):
        t = self.torch
        self.m = len(xyz)
        with t.cuda.stream(self.stream):
            self.q = t.from_numpy(np.ascontiguousarray(xyz).view(np.uint8).reshape(-1).copy()).to(self.dev, non_blocking=False)
            self.local = t.empty(max(self.m, 1), dtype=t.int64, device=self.dev)
            self.hit = t.empty(max(self.m, 1) * 4, dtype=t.float32, device=self.dev)
            self.acc = t.zeros(32, dtype=t.float64, device=self.dev)

    def search(self, pose7, max_dist=0.3):
        p = np.ascontiguousarray(pose7, np.float32)
        self.lo._check(self.ctx.lib.lo_shard_search_dev(self.grid.h, C.c_void_p(self.q.data_ptr()), self.m, p.ctypes.data_as(C.c_void_p),
                                                        C.c_float(max_dist), C.c_void_p(self.local.data_ptr()), C.c_void_p(self.hit.data_ptr())))
        with self.torch.cuda.stream(self.stream):
            return self.local ^ int(SIGN)        # unsigned order -> signed order, so MIN over int64 is MIN over the u64 keys

    def evaluate(self, keys_global, x7d):
        x = np.ascontiguousarray(x7d, np.float64)
        with self.torch.cuda.stream(self.stream):
            glob = keys_global ^ int(SIGN)
        self.lo._check(self.ctx.lib.lo_shard_evaluate_dev(self.grid.h, C.c_void_p(self.q.data_ptr()), self.m, C.c_void_p(self.local.data_ptr()),
                                                          C.c_void_p(glob.data_ptr()), C.c_void_p(self.hit.data_ptr()),
                                                          x.ctypes.data_as(C.c_void_p), C.c_void_p(self.acc.data_ptr())))
        self._keep = glob                       # alive until the kernels have run
        return self.acc


class TorchCollective:
    """Reductions over the local shards, then over torch.distributed (if initialised)."""

    def __init__(self, dist=None):
        self.dist = dist if dist is not None and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 else None

    def min_keys(self, tensors, stream=None):
        import torch
        ctxm = torch.cuda.stream(stream) if stream is not None else _null()
        with ctxm:
            out = tensors[0].clone()
            for t in tensors[1:]:
                out = torch.minimum(out, t.to(out.device))
            if self.dist is not None:
                self.dist.all_reduce(out, op=self.dist.ReduceOp.MIN)
        return out

    def sum_acc(self, tensors, stream=None):
        import torch
        ctxm = torch.cuda.stream(stream) if stream is not None else _null()
        with ctxm:
            out = tensors[0].clone()
            for t in tensors[1:]:
                out = out + t.to(out.device)
            if self.dist is not None:
                self.dist.all_reduce(out, op=self.dist.ReduceOp.SUM)
            return out.cpu().numpy()            # 256 B to the host: the LM step logic runs there


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class HostLM:
    """ceres::Solve's step logic through the C ABI (lo_lm_*, csrc/lo_lm.cuh compiled for the host)."""

    def __init__(self, lib):
        self.lib = lib
        self.h = C.c_void_p(lib.lo_lm_create())

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.lo_lm_destroy(self.h)
            self.h = None

    def begin(self, x0, acc32, prior_b):
        x0, acc32, prior_b = (np.ascontiguousarray(a, np.float64) for a in (x0, acc32, prior_b))
        return self.lib.lo_lm_begin(self.h, x0.ctypes.data_as(C.c_void_p), acc32.ctypes.data_as(C.c_void_p), prior_b.ctypes.data_as(C.c_void_p)) == 1

    def propose(self):
        cand = np.zeros(7)
        need = self.lib.lo_lm_propose(self.h, cand.ctypes.data_as(C.c_void_p)) == 1
        return need, cand

    def update(self, acc32):
        acc32 = np.ascontiguousarray(acc32, np.float64)
        return self.lib.lo_lm_update(self.h, acc32.ctypes.data_as(C.c_void_p)) == 1

    def result(self):
        x, step, cost = np.zeros(7), C.c_double(0), C.c_double(0)
        term, rec = C.c_int32(0), C.c_int32(0)
        self.lib.lo_lm_result(self.h, x.ctypes.data_as(C.c_void_p), C.byref(step), C.byref(cost), C.byref(term), C.byref(rec))
        return x, step.value, cost.value, term.value, rec.value


class ShardedCloudMatcher:
    """align() over `shards` (this process's shards) and `coll` (cross-process reductions)."""

    def __init__(self, shards, coll, lm, stream=None):
        self.shards, self.coll, self.lm, self.stream = shards, coll, lm, stream
        self.evaluations = 0

    def _evaluate(self, keys, x):
        self.evaluations += 1
        return self.coll.sum_acc([s.evaluate(keys, x) for s in self.shards], self.stream)

    def align(self, guess7, max_dist=0.3):
        pose = np.ascontiguousarray(guess7, np.float32).copy()
        prior_b = pose[:3].astype(np.float64)
        log = []
        for outer in range(35):                                                  # cloud_matcher.cpp:117
            x0 = np.array([pose[3], pose[4], pose[5], pose[6], pose[0], pose[1], pose[2]], np.float64)   # :122-131
            keys = self.coll.min_keys([s.search(pose, max_dist) for s in self.shards], self.stream)       # exchange 1
            acc = self._evaluate(keys, x0)                                                                # exchange 2
            cont = self.lm.begin(x0, acc, prior_b)
            pairs = int(acc[28] + 0.5)
            while True:                                                          # ceres::Solve, <= 4 iterations
                need, cand = self.lm.propose() if cont else (False, None)
                if not need:
                    break
                cont = self.lm.update(self._evaluate(keys, cand))
            x, step_norm, cost, term, rec = self.lm.result()
            pose = np.array([x[4], x[5], x[6], x[0], x[1], x[2], x[3]]).astype(np.float32)                # :161-167
            log.append((pairs, rec, step_norm, cost, term))
            if step_norm < 1e-4 and outer > 3:                                   # :169
                break
        w, x_, y, z = pose[3], pose[4], pose[5], pose[6]                          # :175 Quaternionf::normalize()
        nrm = np.sqrt(np.float32((x_ * x_ + z * z) + (y * y + w * w)))
        pose[3:] = np.array([w / nrm, x_ / nrm, y / nrm, z / nrm], np.float32)
        return pose, log
