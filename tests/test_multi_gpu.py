"""Real multi-GPU runs of the sharded Plan() protocols (NCCL and peer-memory transports), launched with torchrun from inside
pytest.  Skipped on boxes with fewer than two GPUs; tests/run_sharded_nccl.py is the script, profiles/ holds a recorded log."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sliced_plan_identical_on_every_rank(world):
    """bench.py --gpus N compares, after its timed loops, every rank's rows and the job-wide counters with the oracle's plan
    of the full snapshot and exits non-zero on a mismatch: the same check the driver's scaling run carries in its line."""
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
                        "--master-port", str(29550 + world), os.path.join(ROOT, "bench.py"), "--gpus", str(world), "--steps", "5", "--warmup", "3",
                        "--workload", "mig-c2"],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    import json
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["parity"] == "identical" and line["n_gpus"] == world


@pytest.mark.parametrize("transport", ["peer-memory", "nccl"])
def test_sharded_protocols_on_two_gpus(transport):
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ)
    env.pop("NOSPLAN_NO_PEER", None)
    if transport == "nccl":
        env["NOSPLAN_NO_PEER"] = "1"
    port = 29541 if transport == "nccl" else 29542
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tests", "run_sharded_nccl.py")],
                       capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MISMATCH" not in r.stdout and f"sliced[{transport}]=OK" in r.stdout
    assert r.stdout.count("sharded-quota=OK") == 2 and "(chained sweep)" in r.stdout
