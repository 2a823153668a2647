"""bench.py's contract: one JSON line on stdout with the keys the driver reads — the CPU reference arm here, the GPU arm on a B200."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
             "data", "config", "e2e"}


def run_bench(*args, timeout=600):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, f"exactly one line on stdout, got {len(lines)}"
    return json.loads(lines[0])


def test_reference_arm_line():
    d = run_bench("--impl", "reference", "--workload", "C1", "--steps", "1", "--warmup", "0")
    assert BASE_KEYS <= set(d), BASE_KEYS - set(d)
    assert d["impl"] == "reference" and d["metric"] == "G iter/s" and d["value"] > 0 and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]
    assert set(d["config"]) == {"workload", "iterations_per_frame"}   # the GPU arm's line carries exactly these two keys as well


@pytest.mark.gpu
def test_gpu_arm_line():
    d = run_bench("--workload", "C1", "--steps", "2", "--warmup", "3", "--no-cpu-baseline", "--no-extras")
    assert BASE_KEYS | {"roofline", "clocks", "gpu_launches", "cpu_baseline"} <= set(d)
    assert d["n_gpus"] == 1 and d["dtype"] == "f64" and d["gpu_launches"] > 0 and d["value"] > 100
    assert set(d["config"]) == {"workload", "iterations_per_frame"} and "kernel" in d["run"]   # same config keys as the reference arm
    assert d["e2e"]["ms_min"] <= d["e2e"]["ms_median"] and d["e2e_fresh_map"]["value"] > 0 and d["e2e_pinned"]["value"] > 0
    r = d["roofline"]
    assert r["bound"] == "fp64" and 0.05 < r["frac"] < 0.84 and r["peak"] > 30 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-3
    assert r["traffic"] is None or r["traffic"] >= 0
    e = d["e2e"]
    assert e["value"] > 0 and e["d2h_bytes_per_step"] >= 1920 * 1080 * 3 and e["value"] < d["value"]
    assert d["clocks"]["sm_mhz"] > 500 and not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert d["extra"]["single_stream"]["ms_per_frame"] > 0
