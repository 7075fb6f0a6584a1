"""CPU: the helper-thread staging copy of lo_odometry_process (csrc/lo_copy_pool.h) — correctness of every byte, wake-up of
sleeping helpers, clean shutdown. Compiled with g++ from the header alone (no CUDA)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_copy_pool(tmp_path):
    exe = str(tmp_path / "copy_pool_test")
    subprocess.run(["g++", "-std=c++17", "-O2", "-Wall", "-pthread", "-o", exe, os.path.join(ROOT, "tests", "csrc", "copy_pool_test.cpp")],
                   check=True, capture_output=True, text=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("ok 3000 copies")
