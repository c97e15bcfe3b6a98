"""bench.py's contract where it can be checked without a GPU: the reference arm runs the full mig-10k configuration on the host
cores and prints one JSON line with the keys the driver reads; our arm refuses to run without a CUDA device (no CPU fallback)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_same_config_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-800:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["metric"] == "planner_candidates_per_sec" and line["unit"] == "candidates/s"
    assert line["higher_is_better"] is True and line["n_gpus"] == 1 and line["gpu_launches"] == 0
    cfg = line["config"]
    assert cfg["workload"] == "mig-10k" and cfg["nodes"] == 10000 and cfg["gpus_per_node"] == 8 and cfg["pods"] == 10000
    assert "sample_nodes" not in cfg and "sample_pods" not in cfg            # the full workload, not a sub-snapshot
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == 1 and cb["value"] == line["value"] and "FULL workload" in cb["sample"]
    # 1e8 candidates per Plan(): the value must be what the time says
    assert abs(line["value"] - 1e8 / (line["ms_per_step"] / 1e3)) / line["value"] < 1e-6


def test_our_arm_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present: nothing to refuse")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode != 0 and not any(l.startswith("{") for l in r.stdout.splitlines())
