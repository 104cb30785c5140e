"""Checks the NVLink peer-memory all-reduce (csrc/b2_peer.cu) against NCCL and times both, under torchrun:

  torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/peer_check.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import torch
import torch.distributed as dist


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    dev = torch.device("cuda", torch.cuda.current_device())
    from upscaler import peer
    red = peer.reducer(dist, dev)
    if red is None:
        if rank == 0:
            print("peer all-reduce unavailable on this box")
        return
    g = torch.Generator(device="cpu").manual_seed(100 + rank)
    worst = 0.0
    for it, n in enumerate([1, 2, 128, 129, 1000, 1024, 2048, 4096] * 8):
        a = torch.randn(n, generator=g).to(dev)
        b = a.clone()
        red.all_reduce(a)
        dist.all_reduce(b)
        worst = max(worst, float((a - b).abs().max()))
        gathered = [torch.empty_like(a) for _ in range(world)]
        dist.all_gather(gathered, a)
        assert all(torch.equal(gathered[0], t) for t in gathered), "replicas disagree bitwise"
    # latency: back-to-back dependent reductions of 128 floats (one BatchNorm's sum and sum of squares for 64 channels)
    res = {}
    x = torch.ones(128, device=dev)
    for name, fn in (("peer", lambda: red.all_reduce(x)), ("nccl", lambda: dist.all_reduce(x))):
        for _ in range(20):
            x.fill_(1.0); fn()
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        big = torch.empty((8192, 8192), device=dev)
        for _ in range(6):                       # ~25 ms of queued GPU work: the host gets ahead, the loop below is GPU-bound
            torch.mm(big, big)
        e0.record()
        for _ in range(200):
            x.mul_(0.5); fn()
        e1.record(); torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / 200 * 1e3
    if rank == 0:
        print("peer all-reduce vs NCCL over %d ranks: max |diff| %.3g, replicas bitwise equal; latency per dependent "
              "reduction (GPU-bound, one tiny kernel in between): peer %.1f us, nccl %.1f us" % (world, worst, res["peer"], res["nccl"]))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
