"""World-size-2 `gloo` tests (CPU) of the host-side data-parallel logic of upscaler/training.py: the bucketed gradient
all-reduce driven by the backward passes' on_ready(offset) hooks, the cross-replica BatchNorm sum, the replica-mean of
the reported losses, and bench.py's per-rank sharding of frames.  The kernels themselves need a GPU (tests -m gpu);
tools/dp_check.py is the 2-GPU NCCL check of the same path."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "video-cycle_gan-upscaling_b200")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from upscaler import training as T
        res = {}
        # ---- bucketed gradient all-reduce: hooks arrive tail-first, small slices merge into the next bucket
        n = 1 << 20
        rng = np.random.default_rng(7 + rank)
        flat = torch.from_numpy(rng.standard_normal(n).astype(np.float32))
        mine = flat.clone()
        red = T.GradReducer(min_elems=1 << 17)
        hook = red.hook(flat, n)
        for off in (n - 1000, n - (1 << 18), n - (1 << 18) - 5, 1 << 19, 1 << 10, 0):
            hook(off)
        red.wait()
        res["buckets"] = list(red.log)
        res["sum"] = flat.numpy().copy()
        res["mine"] = mine.numpy()
        # ---- cross-replica BatchNorm statistics
        sync = T._make_sync()
        st = torch.full((8,), float(rank + 1))
        res["rows"] = sync(st, 10)
        res["st"] = st.numpy().copy()
        # ---- replica mean of the reported losses
        c = T.Compiled("gen", None, None)
        res["loss"] = c._mean_over_ranks(torch.tensor([1.0 + rank, 10.0 * (rank + 1)]))
        # ---- switches: per-replica BatchNorm statistics hand the layers no sync object; CPU tensors never reach a CUDA graph
        T.SYNC_BATCHNORM = False
        res["nosync"] = T._make_sync() is None
        T.SYNC_BATCHNORM = True
        res["sync_again"] = T._make_sync() is not None
        res["graph_cpu"] = c._graph_lookup(("gen", (1,), (1,), None, world, ()), torch.device("cpu"), lambda s_, p_: 1 / 0) is None
        # ---- decisions that change the collectives of a step are taken together: a range hit on rank 0 ALONE moves both
        # replicas to bf16 storage (one rank switching by itself would change its graph key and all-reduce pattern), and a
        # capture that fails on one rank keeps every rank on the eager path
        import warnings
        from upscaler.networks import UpscalerOrigNet
        net = UpscalerOrigNet(kernel_size=3, device="cpu")
        res["dtype_before"] = net.train_dtype
        if rank == 0:
            net.range_flag().fill_(1.0)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            c._range_tick(net)                                  # first tick always reads the flag
        res["dtype_after"] = net.train_dtype
        c._range_tick(net)                                      # bf16 now: no collective, no read
        net2 = UpscalerOrigNet(kernel_size=3, device="cpu")
        c._range_tick(net2)                                     # nobody hit: nobody switches
        res["dtype_quiet"] = net2.train_dtype
        res["any"] = [T.agree(rank == 1, None, "any"), T.agree(False, None, "any")]
        res["all"] = [T.agree(rank == 1, None, "all"), T.agree(True, None, "all")]
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def test_dp_host_logic_world2_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n = 1 << 20
    total = got[0]["mine"] + got[1]["mine"]
    for r in range(world):
        np.testing.assert_allclose(got[r]["sum"], total, rtol=0, atol=0)       # every element reduced exactly once
        b = got[r]["buckets"]
        assert b[0][1] == n and b[-1][0] == 0
        assert all(b[i][0] == b[i + 1][1] for i in range(len(b) - 1))           # contiguous, tail first
        assert all(hi - lo >= (1 << 17) for lo, hi in b[:-1])                   # small slices were merged
        assert got[r]["rows"] == 20
        np.testing.assert_allclose(got[r]["st"], 3.0)
        assert got[r]["loss"] == [1.5, 15.0]
        assert got[r]["nosync"] and got[r]["sync_again"] and got[r]["graph_cpu"]
        assert got[r]["dtype_before"] == torch.float16 and got[r]["dtype_after"] == torch.bfloat16
        assert got[r]["dtype_quiet"] == torch.float16
        assert got[r]["any"] == [True, False] and got[r]["all"] == [False, True]
    assert got[0]["buckets"] == got[1]["buckets"]


def test_gradreducer_single_process_is_a_no_op():
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    from upscaler import training as T
    flat = torch.arange(1000, dtype=torch.float32)
    red = T.GradReducer(min_elems=100)
    hook = red.hook(flat, 1000)
    for off in (990, 700, 650, 0):
        hook(off)
    red.wait()
    assert red.log == [(700, 1000), (0, 700)]
    assert torch.equal(flat, torch.arange(1000, dtype=torch.float32))
    assert T._make_sync() is None


def test_generator_backward_hook_offsets_cover_the_flat_buffer():
    """The offsets UpscalerOrigNet.backward reports are layer starts of the flat layout, in descending order."""
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    from upscaler.networks import UpscalerOrigNet
    net = UpscalerOrigNet(kernel_size=3, device="cpu")
    offs = [net.offsets[("upscaling/0/block/conv_transp", 0)][0], net.offsets[("res_block/8/conv_pre", 0)][0], 0]
    assert offs == sorted(offs, reverse=True) and offs[0] < net.n_trainable
    # everything behind an offset belongs to layers that backward has finished by then
    tail_layers = {name for (name, j), (o, n, shp) in net.offsets.items() if o >= offs[0] and o < net.n_trainable}
    assert tail_layers == {"upscaling/0/block/conv_transp", "upscaling/1/block/conv_transp", "final/conv"}
