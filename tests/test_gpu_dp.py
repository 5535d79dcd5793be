"""Data-parallel gradient parity on hardware: 2 NCCL ranks, each with a shard of the batch, against the
average of the per-shard gradients computed on one GPU - the reference's sketched tower averaging
(src/deepSpeech_train.py:193-228), batch-norm statistics per tower.  fp32 path to fp32 round-off (5e-5; 6e-6 typical), bf16 tensor-core
path to 2e-2 (its split-K reductions are not bit-reproducible).  Skipped on a box with fewer than 2 GPUs."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("path", ["fp32", "bf16"])
def test_two_rank_nccl_gradients_match_tower_average(path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "dp_nccl_worker.py"), path]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "DP_PARITY_OK" in r.stdout, r.stdout[-2000:]
