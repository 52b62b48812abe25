"""bench.py's one-line JSON contract: the reference arm on CPU (here), the fsx arm on a GPU."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
             "data", "config", "e2e", "cpu_baseline"}


def _run(*args):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    return json.loads(out.stdout.strip().splitlines()[-1])


def test_reference_arm_prints_the_contract_line():
    d = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-rows", "1000000")
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["metric"] == "groupby_agg_rows_per_s" and d["unit"] == "rows/s" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_stdout_carries_the_json_line_and_nothing_else():
    """native libraries write to descriptor 1 as well (NCCL's version banner at communicator creation): bench.py points
    that descriptor at stderr and writes its line to a duplicate of the original one"""
    code = ("import os, sys; sys.path.insert(0, %r); import bench; bench.claim_stdout(); os.write(1, b'NCCL version 0.0.0\\n'); "
            "print('python chatter'); bench.emit({'metric': 'm', 'value': 1.5})") % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout == json.dumps({"metric": "m", "value": 1.5}) + "\n"
    assert "NCCL version 0.0.0" in out.stderr and "python chatter" in out.stderr


@pytest.mark.gpu
def test_fsx_arm_prints_the_contract_line():
    d = _run("--steps", "5", "--warmup", "3", "--no-stats", "--no-c4", "--no-c5", "--no-hj", "--cpu-rows", "5000000")
    assert d["verified"] is True and d["verification"]["checks"]["count_sum"] and d["verification"]["sample_keys"] == 10000
    assert BASE_KEYS | {"gpu_launches", "roofline", "clocks"} <= set(d)
    assert d["n_gpus"] == 1 and d["steps"] == 5 and d["warmup"] == 3 and d["gpu_launches"] > 0 and d["dtype"] == "f64"
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-12 and r["traffic"] > 1.2e9
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 1_200_000_000 and e["d2h_bytes_per_step"] > 0 and 0 < e["value"] < d["value"]
    assert d["cpu_baseline"]["kind"] == "port" and {"sm_mhz", "sm_max_mhz", "reasons"} <= set(d["clocks"])
