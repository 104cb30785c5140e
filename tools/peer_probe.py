"""What does ONE cross-replica BatchNorm exchange cost?  Under torchrun with N ranks:
  * in-kernel: `iters` dependent exchanges of n floats inside a single launch (csrc/b2_peer.cu peer_probe_kernel) -> pure
    exchange latency over NVLink / NVSwitch, no launches, no host;
  * the host side of a training step on every rank: time to enqueue the step vs time blocked in the loss read
    (training.HOST_STATS): a rank whose wait is ~0 is bound by its host's launch rate."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np
import torch
import torch.distributed as dist


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    if os.environ.get("OMP_NUM_THREADS") == "1":
        torch.set_num_threads(max(1, min(16, (os.cpu_count() or 1) // world)))
    from upscaler import peer, training
    red = peer.reducer(dist, dev)
    try:
        clk = int(torch.cuda.clock_rate())           # current SM clock in MHz (NVML)
    except Exception:  # noqa: BLE001
        clk = 1965
    if red is not None:
        for n in (128, 2048):
            red.exchange_probe(n, 50)
            dist.barrier()
            cyc = red.exchange_probe(n, 2000)
            if rank == 0:
                print("in-kernel exchange of %d floats over %d ranks: %.0f SM cycles = %.2f us at %d MHz" % (n, world, cyc, cyc / clk, clk))
    from upscaler.model import VGG_MSE_LOSS, compile_training_model, make_upscaler_orig
    up = make_upscaler_orig((384, 384, 3), kernel_size=3)
    tm = compile_training_model(up, VGG_MSE_LOSS((384, 384, 3), 1e-3).loss)
    rng = np.random.default_rng(rank)
    for B in (16, 2):
        hr = rng.uniform(-1, 1, (B, 384, 384, 3)).astype(np.float32)
        x = hr.reshape(B, 96, 4, 96, 4, 3).mean(axis=(2, 4)).astype(np.float32)
        for _ in range(3):
            tm.train_on_batch(x, hr)
        dist.barrier(); torch.cuda.synchronize()
        training.HOST_STATS.update(wait_s=0.0, reads=0)
        steps = 8
        t0 = time.perf_counter()
        for _ in range(steps):
            tm.train_on_batch(x, hr)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / steps
        wait = training.HOST_STATS["wait_s"] / steps
        t = torch.tensor([dt * 1e3, wait * 1e3, (dt - wait) * 1e3], device=dev)
        tmax, tmin = t.clone(), t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        if rank == 0:
            print("train step B=%d per GPU, %d ranks (host cores %d, torch threads %d): step %.2f ms; host blocked in the loss read "
                  "min %.2f / max %.2f ms over the ranks; host enqueue time min %.2f / max %.2f ms" % (
                      B, world, os.cpu_count() or 0, torch.get_num_threads(), float(tmax[0]), float(tmin[1]), float(tmax[1]),
                      float(tmin[2]), float(tmax[2])))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
