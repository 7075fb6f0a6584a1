"""Host-pointer path timing: lo_odometry_process on pinned scans, wall clock and CUDA events, no L2 flush."""
import ctypes as C, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import lidar_odometry_demo_b200 as lo
from bench import make_stream
n = int(sys.argv[1]) if len(sys.argv) > 1 else 120
scans = make_stream(n, 0)
ctx = lo.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", 0))
pinned = [torch.from_numpy(s.view(np.uint8).reshape(-1).copy()).pin_memory() for s in scans]
dev = [p.cuda() for p in pinned]
torch.cuda.synchronize()
for name, call in (("host", lambda od, k: ctx.lib.lo_odometry_process(od.h, C.c_void_p(pinned[k].data_ptr()), len(scans[k]))),
                   ("resident", lambda od, k: ctx.lib.lo_odometry_process_dev(od.h, C.c_void_p(dev[k].data_ptr()), len(scans[k])))):
    od = lo.LidarOdometry(ctx)
    for k in range(10):
        call(od, k)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(10, n):
        assert call(od, k) == 0
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"{name}: {1e3 * dt / (n - 10):.4f} ms/scan wall, {(n - 10) / dt:.0f} scans/s")
