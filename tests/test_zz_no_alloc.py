"""GPU: lo_nn_residual_static_dev — the fused NN + residual pass for callers that cannot allocate (CUDA graph capture,
serving loops). It has no counterpart in the reference: same kernels as
lo_nn_residual_dev, with the per-block partial sums in a fixed device array inside the library and the result in a device
block the caller owns. Checked through the C ABI against the existing entry point on two calls in a row (the array is
shared by every call and keeps the last call's contents), then on the size limit.

Collected last on purpose. Self-contained: it builds its own context and grid, so it
passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

import lidar_odometry_demo_b200 as lo
from lidar_odometry_demo_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def _eq(e):
    return (float(e.cost), np.array(e.g[:]), np.array(e.jtj[:]), int(e.n_valid), int(e.candidates), int(e.occupied))


def test_static_scratch_entry_point_matches_the_allocating_one():
    import torch   # test plumbing only: device buffers for the caller-owned input and output
    ctx = lo.Context(0)
    lib = ctx.lib
    cloud = S.planar_patch_map(200_000, seed=5, extent=(60.0, 60.0, 12.0))
    g = lo.VoxelGrid(ctx, 0.2, 20, capacity_hint=len(cloud) // 8)
    g.addCloud(cloud)
    rng = np.random.default_rng(3)
    dev = torch.device("cuda", 0)
    d_out = torch.zeros(C.sizeof(lo.NormalEq), dtype=torch.uint8, device=dev)   # the caller's output block
    b = np.zeros(3)
    # two calls in a row, different sizes (different numbers of blocks write the shared array) and poses
    for m, pose in ((40_000, [0.02, -0.01, 0.015, 0.9999875, 0.0, 0.0, 0.005]), (3_000, [-0.03, 0.02, 0.0, 0.99998, 0.002, -0.006, 0.0])):
        sel = rng.choice(len(cloud), m, replace=False)
        q = np.zeros(m, S.XYZ)
        q["x"], q["y"], q["z"] = cloud["x"][sel] + 0.03, cloud["y"][sel] - 0.02, cloud["z"][sel] + 0.01
        far = rng.random(m) < 0.2                      # some queries without a correspondence
        q["z"][far] += np.float32(0.5)
        d_q = torch.from_numpy(q.view(np.uint8).reshape(-1)).to(dev)
        pose = np.array(pose, np.float32)
        x = np.array([pose[3], pose[4], pose[5], pose[6], pose[0], pose[1], pose[2]], np.float64)
        args = (g.h, C.c_void_p(d_q.data_ptr()), m, pose.ctypes.data_as(C.c_void_p), x.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p),
                C.c_float(0.3))
        want = lo.NormalEq()
        assert lib.lo_nn_residual_dev(*args, C.byref(want)) == 0             # existing entry point: host result, synchronises
        assert lib.lo_nn_residual_static_dev(*args, C.c_void_p(d_out.data_ptr())) == 0, lib.lo_last_error()
        ctx.synchronize()                                                   # the new one only enqueues
        got = lo.NormalEq.from_buffer_copy(d_out.cpu().numpy().tobytes())
        cw, gw, hw, nw, candw, occw = _eq(want)
        cg, gg, hg, ng, candg, occg = _eq(got)
        assert 0 < nw < m and (ng, candg, occg) == (nw, candw, occw)        # same correspondences, same visits
        # the tolerance of tests/test_gpu_tile_search.py (the two paths run the same kernels in the same launch shape)
        assert abs(cg - cw) <= 1e-12 * abs(cw)
        assert np.allclose(gg, gw, rtol=1e-10, atol=1e-9) and np.allclose(hg, hw, rtol=1e-10, atol=1e-9)
    # sizes beyond the stated maximum are refused with a clear error, nothing is launched
    launches = ctx.kernel_launches
    rc = lib.lo_nn_residual_static_dev(g.h, C.c_void_p(d_q.data_ptr()), (1 << 22) + 1, pose.ctypes.data_as(C.c_void_p),
                                       x.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), C.c_float(0.3), C.c_void_p(d_out.data_ptr()))
    assert rc == lo.LO_E_CAPACITY and b"at most" in lib.lo_last_error() and ctx.kernel_launches == launches
    g.close()
    ctx.close()
