import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    """The CPU oracle (self-contained restatement, oracle/lo_oracle.cpp)."""
    import oracle
    return oracle.load("port")


@pytest.fixture(scope="session")
def ref():
    """The reference's own headers compiled into oracle/_ref (present where /root/reference was mounted at build time)."""
    import oracle
    if not oracle.available("reference"):
        if os.path.isdir("/root/reference/src"):
            oracle.build("reference")
        else:
            pytest.skip("oracle/_ref/liblo_ref.so not built (needs /root/reference)")
    return oracle.load("reference")


@pytest.fixture(scope="session")
def ctx():
    import lidar_odometry_demo_b200 as lo
    c = lo.Context(0)  # raises without a CUDA device: there is no CPU fallback
    yield c
    c.close()


@pytest.fixture(scope="session")
def stream12():
    """12 synthetic VLP16 scans of BASELINE.json config 2 (seeds of SURVEY.md section 8(d))."""
    from lidar_odometry_demo_b200 import synthetic as S
    scans, pos, q = S.vlp16_stream(12)
    return scans, pos, q


def fields_equal(a, b, names):
    return all(np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)) for k in names)
