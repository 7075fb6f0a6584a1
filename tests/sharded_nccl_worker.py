"""One rank of the native sharded matcher (lo_sharded_*): `python sharded_nccl_worker.py <rank> <world> <dir>`.
Rendezvous through files in <dir>: rank 0 writes the NCCL unique id, every rank loads the same scene (scene.npz), builds
its shard from the replicated key-frame cloud, aligns the replicated queries and writes its pose + stats."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lidar_odometry_demo_b200 as lo  # noqa: E402


def main():
    rank, world, d = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
    ctx = lo.Context(rank)
    id_path = os.path.join(d, "nccl_id.bin")
    if rank == 0:
        uid = lo.sharded_unique_id()
        with open(id_path + ".tmp", "wb") as f:
            f.write(uid.tobytes())
        os.replace(id_path + ".tmp", id_path)
    else:
        t0 = time.time()
        while not os.path.exists(id_path):
            if time.time() - t0 > 120:
                raise RuntimeError("no NCCL id from rank 0")
            time.sleep(0.05)
        uid = np.fromfile(id_path, np.uint8)
    z = np.load(os.path.join(d, "scene.npz"))
    kf, q, guess = z["kf"], z["q"], z["guess"]
    m = lo.ShardedCloudMatcher(ctx, uid, rank, world, 0.2, 20)
    m.grid.addCloud(kf)
    out = {"rank": rank, "voxels": m.grid.size(), "points": m.grid.numPoints(), "runs": []}
    for rep in range(3):                                   # repeated aligns: the mailbox sequence keeps counting
        t0 = time.perf_counter()
        pose, stats = m.align(q, guess, with_stats=True)
        dt = time.perf_counter() - t0
        out["runs"].append({"pose": [float(v) for v in pose], "pose_bits": pose.view(np.uint32).tolist(), "outer": stats.outer_iterations,
                            "pairs": list(stats.pairs_log[:stats.outer_iterations]), "cost": stats.final_cost, "seconds": dt})
    empty, _ = m.align(q[:0], guess, with_stats=True)       # no queries: prior-only solve, no exchange
    out["empty_pose"] = [float(v) for v in empty]
    info = m.info()
    out["info"] = {k: int(getattr(info, k)) for k, _ in info._fields_}
    with open(os.path.join(d, f"result{rank}.json"), "w") as f:
        json.dump(out, f)
    m.close()
    ctx.close()


if __name__ == "__main__":
    main()
