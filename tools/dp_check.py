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


def run(up, tm, lo, hi, steps, B, reverse=False, record=None, replay=None):
    """Returns the losses of `steps` steps and the (all-reduced) flat gradient of the FIRST step.  Adam's first updates are
    ~lr*sign(g) whatever the gradient's size, so two free-running trajectories drift apart chaotically (0.6-2 % of the loss
    after three steps, run to run); the single-GPU run therefore REPLAYS the data-parallel run's weights before every step
    (`record` / `replay`: snapshots of the flat parameter buffer), which makes every step's loss a like-for-like comparison
    of N shards against one batch."""
    losses, g0 = [], None
    for s in range(steps):
        x, hr = data(B, s)
        if record is not None:
            record.append(up.net.flat.detach().clone())
        if replay is not None:
            up.net.flat.copy_(replay[s])
            up.net._train_dirty = up.net._infer_dirty = True
        if reverse:      # same batch, samples in reverse order: identical mathematics, different tiles and summation order
            x, hr = x[::-1].copy(), hr[::-1].copy()
        losses.append(tm.train_on_batch(x[lo:hi], hr[lo:hi]))
        if s == 0:
            torch.cuda.synchronize()
            # fp16 training storage carries the gradients multiplied by the step's scale (training.GRAD_SCALE: the number of
            # generated elements of THIS rank's shard); divide it out so shards of different sizes compare
            S = tm._compiled._scale_for(up.net, hr[lo:hi].size, hi - lo)
            g0 = up.net.grad_flat.detach().cpu().numpy().astype(np.float64) / S
    torch.cuda.synchronize()
    return np.array(losses), g0


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
    snaps = []
    l_dp, w_dp = run(up, tm, rank * per, (rank + 1) * per, steps, B, record=snaps)
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
    l_1, w_1 = run(up1, tm1, 0, B, steps, B, replay=snaps)
    rel_l = np.abs(l_dp - l_1) / np.abs(l_1)
    # Round-off floor: the same single-GPU step with the batch in reverse order (identical mathematics).  Gradients travel
    # in 16-bit storage through up to 37 layers, so two mathematically equal runs differ by a few per cent per tensor; the
    # data-parallel run must agree with the single-GPU run as well as the single-GPU run agrees with itself.
    up2, tm2 = build()
    _, g_rev = run(up2, tm2, 0, B, 1, B, reverse=True)
    bad, worst_cos, worst = [], 1.0, (0.0, 0.0, "")
    net = up1.net
    for (name, j), (o, n, shp) in net.offsets.items():
        if o >= net.n_trainable:
            continue
        a, b, c = (v[o:o + n].astype(np.float64) for v in (w_dp, w_1, g_rev))
        na, nb = float(np.linalg.norm(a)), float(np.linalg.norm(b))
        if nb < 1e-6 * np.sqrt(n):        # analytically-zero gradients (conv bias in front of a train-mode BatchNorm)
            continue
        cos, ratio = float(np.vdot(a, b) / (na * nb)), na / nb
        d_dp, d_floor = float(np.linalg.norm(a - b)) / nb, float(np.linalg.norm(c - b)) / nb
        worst_cos = min(worst_cos, cos)
        if d_dp > worst[0]:
            worst = (d_dp, d_floor, "%s[%d]" % (name, j))
        # norm ratio: 6 %, or half the tensor's own round-off floor where that is larger (the 64-element bias gradients of
        # initial/conv sit at a 17-19 % reversed-batch floor)
        if cos < 0.97 or abs(ratio - 1) > max(0.06, 0.5 * d_floor) or d_dp > 2.0 * d_floor + 1e-2:
            bad.append((name, j, round(cos, 4), round(ratio, 4), round(d_dp, 4), round(d_floor, 4)))
    print("DPCHECK world=%d losses dp %s single %s" % (world, np.round(l_dp, 6), np.round(l_1, 6)))
    print("DPCHECK first-step loss rel diff %.2e, max rel loss diff over %d steps %.2e, first-step gradients: worst cosine "
          "%.4f, worst per-tensor rel L2 diff %.3e (same tensor, reversed-batch floor %.3e: %s), out-of-tolerance %s, replica "
          "drift %.2e, buckets %s" % (rel_l[0], steps, rel_l.max(), worst_cos, worst[0], worst[1], worst[2], bad, drift, buckets))
    # every step starts from the same weights on both sides: the losses must agree like the first one does
    ok = rel_l.max() < 1e-3 and not bad and drift == 0.0
    print("DPCHECK", "OK" if ok else "MISMATCH")
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
