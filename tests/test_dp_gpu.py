"""-m gpu: data-parallel training over NCCL on 2 GPUs reproduces the single-GPU step (tools/dp_check.py).  Skipped on a
one-GPU box (the round-end GPU tier); run with `gpurun --gpus 2 -- python -m pytest tests/test_dp_gpu.py -m gpu`."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_replicas_equal_one_big_batch():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29511", os.path.join(ROOT, "tools", "dp_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    print(res.stdout[-3000:], res.stderr[-3000:])
    assert res.returncode == 0
    assert "DPCHECK OK" in res.stdout
