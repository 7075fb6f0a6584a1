import sys, os, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import lidar_odometry_demo_b200 as lo, oracle
from lidar_odometry_demo_b200 import synthetic as S
from util import planar_scene
port = oracle.load("port")
scans, _, _ = S.vlp16_stream(2)
kf = planar_scene(port, scans, 0)
ds = port.grid(0.3, 1); ds.add_cloud(planar_scene(port, scans, 1)); q = ds.get_cloud_xyz()
G = np.array([0.1, 0.05, -0.02, 0.9999619, 0, 0, 0.0087265], np.float32)
ctx = lo.Context(0)
g = lo.VoxelGrid(ctx, 0.2, 20); g.addCloud(kf)
print("single", lo.CloudMatcher().align(g, q, G))
for world in (1, 2):
    ctxs = [lo.Context(0) for _ in range(world)]
    grp = lo.ShardedGroup(ctxs, 0.2, 20); grp.addCloud(kf)
    for m in (0, 64, len(q)):
        try:
            poses, st = grp.align(q[:m], G)
            print("world", world, "m", m, "ok", poses[0], st.outer_iterations)
        except lo.LoError as e:
            print("world", world, "m", m, "ERR", e)
    grp.close()
