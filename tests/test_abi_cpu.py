"""CPU suite: the C-ABI library loads without a GPU and exports every symbol include/lo_b200.h declares;
host-only entry points (params) behave like the reference's Params (src/lidar_odometry.h:23-63)."""
import ctypes as C
import os
import re

import pytest

import lidar_odometry_demo_b200 as lo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "lo_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lo_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    names = declared_symbols()
    assert len(names) >= 35
    assert sorted(lo.SIGNATURES) == names


def test_library_exports_every_declared_symbol():
    lib = lo.load_library()
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_params_defaults_and_yaml(tmp_path):
    lib = lo.load_library()
    p = lo.Params()
    lib.lo_params_default(C.byref(p))
    # GetROSDeclaration defaults (src/lidar_odometry.h:36-48) == config/params.yaml:4-11
    assert (round(p.lidar_min_range, 6), round(p.lidar_max_range, 6)) == (4.0, 80.0)
    assert round(p.keyframe_voxel_size, 6) == 0.2 and p.keyframe_max_points_cnt == 20
    assert round(p.keyframe_matching_voxel_size, 6) == 0.3 and round(p.keyframe_update_voxel_size, 6) == 0.1
    assert round(p.keyframe_cleanup_range, 6) == 80.0 and round(p.angular_divergence_threshold, 6) == 5.0
    y = tmp_path / "params.yaml"
    y.write_text("/**:\n  ros__parameters:\n    publish_tf: true\n    lidar_odometry:\n      lidar_min_range: 2.5\n"
                 "      lidar_max_range: 120.0\n      keyframe_voxel_size: 0.1\n      keyframe_max_points_cnt: 12\n"
                 "      keyframe_matching_voxel_size: 0.05   # comment\n      keyframe_update_voxel_size: 0.05\n"
                 "      keyframe_cleanup_range: 200.0\n      angular_divergence_threshold: 7.5\n")
    q = lo.params_from_yaml(str(y))
    assert (q.lidar_min_range, q.lidar_max_range, q.keyframe_max_points_cnt) == (2.5, 120.0, 12)
    assert round(q.keyframe_voxel_size, 6) == 0.1 and round(q.keyframe_matching_voxel_size, 6) == 0.05
    assert q.keyframe_cleanup_range == 200.0 and q.angular_divergence_threshold == 7.5
    with pytest.raises(lo.LoError):
        lo.params_from_yaml(str(tmp_path / "missing.yaml"))


def test_no_cpu_fallback():
    """Without a CUDA device the context cannot be created: the product never routes through the oracle."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(lo.LoError):
        lo.Context(0)
    src = open(os.path.join(ROOT, "lidar_odometry_demo_b200", "__init__.py")).read()
    assert "import oracle" not in src and "from oracle" not in src
