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


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_peer_allreduce_matches_nccl_on_two_gpus():
    """tools/peer_check.py: the NVLink peer-memory all-reduce (csrc/b2_peer.cu) equals NCCL's and all replicas hold the
    same bits, for 1..4096 elements over 64 consecutive epochs."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29512", os.path.join(ROOT, "tools", "peer_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    print(res.stdout[-2000:], res.stderr[-2000:])
    assert res.returncode == 0 and "replicas bitwise equal" in res.stdout


def test_peer_allreduce_single_rank_and_errors():
    """world = 1 (the region is the process's own): the reduction is the identity through both parities of the staging
    buffer; bad arguments are reported through b2_last_error, not launched."""
    import ctypes
    sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
    from upscaler import _lib
    lib = _lib.load()
    own, handle = ctypes.c_void_p(), ctypes.create_string_buffer(64)
    _lib.check(lib.b2_peer_alloc(ctypes.byref(own), handle), "b2_peer_alloc")
    regions = (ctypes.c_void_p * 1)(own.value)
    nmax = lib.b2_peer_max_elems()
    assert nmax >= 2048                    # the widest BatchNorm on the path: 1024 channels x (sum, sum of squares)
    for epoch, n in enumerate([1, 31, 128, 1000, nmax, 64, 64, 64], start=1):
        x = torch.randn(n, device="cuda")
        ref = x.clone()
        if epoch % 2:       # plain launch: the epoch travels as an argument
            rc = lib.b2_peer_allreduce_f32(_lib.ptr(x), n, regions, 0, 1, epoch, None, _lib.stream_ptr())
        else:               # graph-replay form: position 1 inside the step + the device-side base (include/b2srgan.h)
            base = torch.full((1,), epoch - 1, dtype=torch.int32, device="cuda")
            rc = lib.b2_peer_allreduce_f32(_lib.ptr(x), n, regions, 0, 1, 1, _lib.ptr(base), _lib.stream_ptr())
        _lib.check(rc, "b2_peer_allreduce_f32")
        torch.cuda.synchronize()
        assert torch.equal(x, ref)
    x = torch.zeros(8, device="cuda")
    assert lib.b2_peer_allreduce_f32(_lib.ptr(x), nmax + 1, regions, 0, 1, 99, None, _lib.stream_ptr()) != 0
    assert b"exceed" in lib.b2_last_error()
    assert lib.b2_peer_allreduce_f32(_lib.ptr(x), 8, regions, 3, 2, 99, None, _lib.stream_ptr()) != 0
    _lib.check(lib.b2_peer_close(own, 0), "b2_peer_close")
