"""Where does the data-parallel step lose time at N ranks?  Under torchrun: the config-3 generator step (16 patches per GPU)
timed four ways -- CUDA-graph replay or eager launches, BatchNorm sums exchanged between the replicas or kept per replica
(training.SYNC_BATCHNORM = False: no exchange on the critical path; not the reference's semantics, a measurement here)."""
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
    from upscaler import training
    from upscaler.model import VGG_MSE_LOSS, compile_training_model, make_upscaler_orig
    rng = np.random.default_rng(rank)
    hr = rng.uniform(-1, 1, (16, 384, 384, 3)).astype(np.float32)
    x = hr.reshape(16, 96, 4, 96, 4, 3).mean(axis=(2, 4)).astype(np.float32)
    for graph in ((True, False) if not os.environ.get('B2_GRAPH_ONLY') else (True,)):
        for syncbn in (True, False):
            training.GRAPH_STEPS, training.SYNC_BATCHNORM = graph, syncbn
            up = make_upscaler_orig((384, 384, 3), kernel_size=3)
            tm = compile_training_model(up, VGG_MSE_LOSS((384, 384, 3), 1e-3).loss)
            for _ in range(6):
                tm.train_on_batch(x, hr)
            dist.barrier(); torch.cuda.synchronize()
            training.HOST_STATS.update(wait_s=0.0, reads=0, stage_wait_s=0.0)
            t0 = time.perf_counter()
            steps = 10
            for _ in range(steps):
                tm.train_on_batch(x, hr)
            torch.cuda.synchronize()
            t = torch.tensor([(time.perf_counter() - t0) / steps * 1e3, training.HOST_STATS.get("stage_wait_s", 0.0) / steps * 1e3,
                              training.HOST_STATS["wait_s"] / steps * 1e3], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if rank == 0:
                print("%d ranks, %s, BatchNorm sums %s: %.2f ms per step (host: waited %.2f ms for its staging threads, %.2f ms in the "
                      "loss read; max over ranks)" % (
                          world, "CUDA-graph replay" if graph else "eager launches", "exchanged (global-batch statistics)" if syncbn else
                          "per replica (no exchange)", float(t[0]), float(t[1]), float(t[2])), flush=True)
            del tm, up
            torch.cuda.empty_cache()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
