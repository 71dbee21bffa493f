"""GPU suite, needs >= 2 devices (skipped on a 1-GPU box): row-sharded training in both collective modes --
NVLink peer memory fused into the scan kernel, and NCCL all-reduce -- must reproduce the oracle's forest bit for bit
at every world size the box offers (2, 4, 8): integer histogram sums make the result independent of the sharding."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_training_matches_oracle(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr",
           "127.0.0.1", "--master-port", str(29541 + world), os.path.join(here, "_mgpu_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    sys.stdout.write(res.stdout[-3000:])
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert res.stdout.count("-> OK") == 6


def test_search_grid_dealt_over_gpus_matches_one_gpu():
    """SURVEY 8f-4, second half: one whole fit per GPU for the tuning step's fold x candidate grid."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    world = min(4, torch.cuda.device_count())
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr",
           "127.0.0.1", "--master-port", "29561", os.path.join(here, "_mgpu_search_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    sys.stdout.write(res.stdout[-2000:])
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "-> OK" in res.stdout
