"""bench.py's reference arm on the CPU: exactly one JSON line on stdout with the keys the driver reads (the arm that times the
oracle port, the only part of bench.py that runs without a GPU), and a loud refusal of the GPU arm when no device is there."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, capture_output=True, text=True, timeout=300)


@pytest.mark.parametrize("workload,metric,unit", [("half_plus_two", "half_plus_two GB/s", "GB/s"),
                                                  ("jpeg", "DecodeJpeg images/sec", "images/s")])
def test_reference_arm_prints_one_json_line(workload, metric, unit):
    p = _run("--impl", "reference", "--workload", workload, "--steps", "1", "--warmup", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == metric and d["unit"] == unit and d["value"] > 0
    assert d["steps"] == 1 and d["warmup"] == 1 and d["higher_is_better"] is True and d["scaling"] == "weak"
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_gpu_arm_refuses_to_run_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = _run("--workload", "half_plus_two", "--steps", "1", "--warmup", "1")
    assert p.returncode != 0
    assert "no CPU path" in (p.stderr + p.stdout)
    assert not [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
