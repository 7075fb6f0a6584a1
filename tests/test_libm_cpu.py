"""CPU suite: csrc/lo_libm.cuh (the device's atan2f / atanf / acosf / sinf) against the host's libm — the functions
the reference calls through Eigen's slerp (src/utils/cloud_transform.h:27) and the classifier's azimuth binning
(src/utils/cloud_classifier.h:49). The sweep program is tests/csrc/libm_sweep.cpp; zero mismatches are required."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_device_libm_matches_host_libm(tmp_path):
    exe = str(tmp_path / "libm_sweep")
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-o", exe, os.path.join(ROOT, "tests", "csrc", "libm_sweep.cpp"), "-lm"], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    rows = {l.split()[0]: (int(l.split()[1]), int(l.split()[2])) for l in out.strip().splitlines()}
    assert set(rows) == {"atan2f", "acosf", "sinf", "atanf", "acosf_near1", "atan2f_special"}
    for name, (bad, total) in rows.items():
        assert total > 0 and bad == 0, (name, bad, total)
