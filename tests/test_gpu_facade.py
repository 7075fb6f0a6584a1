"""GPU: the C++ host facade (host/lo_b200.hpp) over the C ABI, exercised by the reference's own gtest cases
(test/test.cpp) re-expressed in tests/csrc/facade_test.cpp, including MatchingTest on intersection00056.pcd."""
import os
import subprocess

import numpy as np
import pytest

from lidar_odometry_demo_b200 import synthetic as S

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_facade_runs_reference_tests(ctx, tmp_path):
    pkg = os.path.join(ROOT, "lidar_odometry_demo_b200")
    exe = str(tmp_path / "facade_test")
    subprocess.run(["g++", "-std=c++20", "-O2", "-o", exe, os.path.join(ROOT, "tests", "csrc", "facade_test.cpp"), f"-L{pkg}",
                    "-l:liblo_b200.so", f"-Wl,-rpath,{pkg}"], check=True)
    xyz3 = np.load(os.path.join(ROOT, "tests", "golden", "pcd_intersection00056.npz"))["xyz"].astype(np.float32)
    normals = ctx.pca_normals(S.to_xyz(xyz3), 0.25)     # K7: the test's pcl::NormalEstimation on the GPU
    xyz3.tofile(tmp_path / "xyz.f32")
    normals.astype(np.float32).tofile(tmp_path / "nrm.f32")
    r = subprocess.run([exe, str(tmp_path / "xyz.f32"), str(len(xyz3)), str(tmp_path / "nrm.f32")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "PASSED" in r.stdout
