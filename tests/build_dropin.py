"""TEST INFRASTRUCTURE. Builds tests/csrc/_dropin/dropin_test: the reference's own src/lidar_odometry.cpp, compiled unmodified
from where it lies under /root/reference, against the GPU facade (tests/csrc/dropin/dropin.hpp force-included) and linked
with liblo_b200.so. The reference tree does not exist on the GPU box, so the binary is built here (by
__graft_entry__.build()) and travels with the snapshot; tests/test_gpu_dropin.py runs it there."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = "/root/reference/src/lidar_odometry.cpp"
OUT_DIR = os.path.join(ROOT, "tests", "csrc", "_dropin")
EXE = os.path.join(OUT_DIR, "dropin_test")


def build():
    if not os.path.exists(REF_SRC):
        return EXE if os.path.exists(EXE) else None
    os.makedirs(OUT_DIR, exist_ok=True)
    pkg = os.path.join(ROOT, "lidar_odometry_demo_b200")
    src = os.path.join(ROOT, "tests", "csrc", "dropin")
    flags = ["-std=c++23", "-O2", "-ffp-contract=off", "-DLO_B200_WITH_EIGEN", "-DLO_B200_WITH_PCL", f"-I{os.path.join(ROOT, 'oracle', 'shim')}",
             "-include", os.path.join(src, "dropin.hpp")]
    cmd = ["g++"] + flags + ["-o", EXE, REF_SRC, os.path.join(src, "dropin_main.cpp"), f"-L{pkg}", "-l:liblo_b200.so",
                             "-Wl,-rpath,$ORIGIN/../../../lidar_odometry_demo_b200"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("drop-in build failed:\n" + r.stdout + r.stderr)
    return EXE


if __name__ == "__main__":
    print(build())
