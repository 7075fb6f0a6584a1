"""CPU suite: bench.py's reference arm (the reference's CPU path timed on the host cores) runs without a GPU and prints the
contract's JSON line; the product arm refuses to run without a CUDA device instead of falling back to the CPU."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "3", "--warmup", "3"], capture_output=True,
                       text=True, cwd=ROOT, timeout=300, env=dict(os.environ, LO_BENCH_NO_STEADY="1"))
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["metric"] == "registered_scans_per_sec" and line["unit"] == "scans/s"
    assert line["n_gpus"] == 1 and line["steps"] == 3 and line["higher_is_better"] is True and line["vs_baseline"] is None
    assert line["value"] > 0 and abs(line["ms_per_step"] * line["value"] - 1e3) < 1e-6 * 1e3
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] == 1
    assert line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in line["config"] and "model" not in line["config"]
    # `config` must be the very object the product arm prints (the driver compares the two arms' configs)
    sys.path.insert(0, ROOT)
    import bench
    assert line["config"] == bench.shared_config(3, 3)


def test_product_arm_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "2", "--warmup", "3"], capture_output=True, text=True, cwd=ROOT,
                       timeout=300)
    assert r.returncode != 0 and not any(l.startswith('{"metric"') for l in r.stdout.splitlines())
