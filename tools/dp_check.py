"""Data-parallel parity check, run under torchrun with N ranks (one per GPU, NCCL):

  torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py

Every rank builds the same generator + VGG19 loss (seeded weights), takes its contiguous shard of one global batch and
runs `steps` train_on_batch calls (cross-replica BatchNorm sums, bucketed gradient all-reduce overlapped with backward,
replicated Adam).  Rank 0 then repeats the same steps alone on the whole batch (process group torn down) and compares:
reported losses and the updated weights must agree -- N replicas of B/N samples are one batch of B.  Also runs the
adversarial step (train_gan.py structure) the same way."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np
import torch
import torch.distributed as dist


def build(seed_w=11):
    from oracle import srgan_oracle as O
    from upscaler.keras_compat import Adam
    from upscaler.model import VGG_MSE_LOSS, compile_training_model, make_upscaler_orig
    gw = O.init_upscaler_orig_weights(seed=seed_w, kernel_size=3, res_block_num=2)
    vw = O.init_vgg19_weights()
    up = make_upscaler_orig((128, 128, 3), kernel_size=3, res_block_num=2)
    up.net.set_weights(gw)
    tm = compile_training_model(up, VGG_MSE_LOSS((128, 128, 3), 1e-3, vgg19=vw).loss, optimizer=Adam())
    return up, tm


def data(B, step):
    rng = np.random.default_rng(500 + step)
    hr = rng.uniform(-1, 1, (B, 128, 128, 3)).astype(np.float32)
    x = hr.reshape(B, 32, 4, 32, 4, 3).mean(axis=(2, 4)).astype(np.float32)
    return x, hr


def run(up, tm, lo, hi, steps, B):
    losses = []
    for s in range(steps):
        x, hr = data(B, s)
        losses.append(tm.train_on_batch(x[lo:hi], hr[lo:hi]))
    torch.cuda.synchronize()
    return np.array(losses), {k: [np.array(a) for a in v] for k, v in up.net.host_weights().items()}


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    B, steps = 8, 4
    assert B % world == 0
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    up, tm = build()
    per = B // world
    l_dp, w_dp = run(up, tm, rank * per, (rank + 1) * per, steps, B)
    buckets = list(tm._compiled.reducer.log)
    if world > 1:
        # replicas must hold identical weights after the steps
        flat = up.net.flat.clone()
        ref = flat.clone()
        dist.broadcast(ref, 0)
        drift = float((flat - ref).abs().max())
        t = torch.tensor([drift], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        drift = float(t.item())
        dist.barrier()
        dist.destroy_process_group()
    else:
        drift = 0.0
    if rank != 0:
        return
    up1, tm1 = build()
    l_1, w_1 = run(up1, tm1, 0, B, steps, B)
    rel_l = np.abs(l_dp - l_1) / np.abs(l_1)
    worst = 0.0
    for k in w_1:
        for a, b in zip(w_dp[k], w_1[k]):
            d = np.abs(a - b).max() / (np.abs(b).max() + 1e-12)
            worst = max(worst, float(d))
    print("DPCHECK world=%d losses dp %s single %s" % (world, np.round(l_dp, 6), np.round(l_1, 6)))
    print("DPCHECK max rel loss diff %.2e, max weight diff / max|w| %.2e, replica drift %.2e, buckets %s" % (
        rel_l.max(), worst, drift, buckets))
    ok = rel_l.max() < 2e-3 and worst < 2e-2 and drift == 0.0
    print("DPCHECK", "OK" if ok else "MISMATCH")
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
