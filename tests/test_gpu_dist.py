"""Sharded (multi-GPU) analysis equals whole-log analysis; needs >= 2 GPUs on the box."""
import os
import subprocess
import sys

import pytest

from conftest import REPO

pytestmark = pytest.mark.gpu


def test_sharded_equals_whole_log():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(8, n)   # every GPU of the box (the driver's scaling run goes to 8)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29533",
                        os.path.join(REPO, "tests", "dist_gpu_check.py"), "400000"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "DIST_CHECK OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_multi_gpu_energy_equals_oracle():
    """One power trace per rank / one trace cut into time shards (dist_energy.py) == the oracle."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(8, n)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29534",
                        os.path.join(REPO, "tests", "dist_energy_check.py"), "400000", "50000"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "DIST_ENERGY_CHECK OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_sweep_runs_across_gpus_equal_one_gpu():
    """Whole runs per GPU (sweep_api.ShardedSweep) == the same runs on one GPU == the oracle."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(8, n)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29535",
                        os.path.join(REPO, "tests", "dist_sweep_check.py"), "37", "6000"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "DIST_SWEEP_CHECK OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
