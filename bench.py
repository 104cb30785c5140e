#!/usr/bin/env python
"""bench.py — headline benchmark: x4 frame upscaling 270p -> 1080p through the SRGAN generator (BASELINE.json
configs[1], k=3 = the shipped-weights architecture), one process per GPU, frames sharded with no collective.

  python bench.py --gpus N --steps K --warmup W            # our CUDA path
  python bench.py --impl reference --gpus N ...            # the reference's CPU path (oracle restatement; the
                                                           #  reference's TensorFlow is unavailable on this box)

A step = one batch of `--batch` synthetic 270x480 frames per GPU pushed through the generator.
  value : frames/s with inputs resident in HBM (device-timed, CUDA events, max over ranks)
  e2e   : frames/s through the public API Model.predict(host ndarray) -> host ndarray (H2D + D2H inside)
  roofline : the dominant kernel (second Conv2DTranspose, 256->256, 50 % of the FLOPs), timed with CUDA events on the
             launching stream inside the timed region, against MEASURED_PEAKS.json
  cpu_baseline : the CPU oracle on the host cores, 1 frame (bounded sample)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))

GEN_FLOPS = {3: 1227.09e9, 5: 2942.77e9}      # per 270x480 frame, SURVEY.md 8(d)
TRAIN_STEP_FLOPS = {3: 9.69e12, 5: 15.55e12}   # config 3 train step, B = 16 (3G + 3V per sample), SURVEY.md 8(d)
CONVT2_FLOPS = {3: 611.5e9, 5: 1698.7e9}      # second Conv2DTranspose 256->256 at 540x960 -> 1080x1920
H_LR, W_LR = 270, 480


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"bf16_tflops": p.get("bf16_tflops", 1590.0), "bf16_tflops_sustained": p.get("bf16_tflops_sustained", 1400.0),
                "hbm_gbs": p.get("hbm_gbs", 6650.0), "source": "measured"}
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0, "source": "fallback"}


def profiled_dram_traffic(kernel_tag="conv_phases2"):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel, read from the newest committed
    `ncu --set full` summary under profiles/ (tools/ncu_raw_pick.py output; captured at B = 8, k = 3).  None when absent."""
    import glob
    import re
    best = None
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*ncu_full*%s*summary.txt" % kernel_tag))):
        vals = {}
        with open(path) as f:
            for line in f:
                m = re.match(r"(dram__bytes_(?:read|write)\.sum)\s+(\w+)\s+\['([0-9.]+)'", line)
                if m:
                    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(m.group(2))
                    if scale:
                        vals[m.group(1)] = float(m.group(3)) * scale
        if len(vals) == 2:
            best = (sum(vals.values()), os.path.basename(path))
    return best


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for name, col in (("hw_slowdown", 4), ("hw_thermal_slowdown", 5), ("sw_thermal_slowdown", 6),
                                  ("sw_power_cap", 7)):
                    if r[col].lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                pass
        sm.sort()
        # median over the samples taken under load (upper half: the region includes idle samples at both ends)
        load = sm[len(sm) // 2:] if sm else []
        med = load[len(load) // 2] if load else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def cpu_oracle_frames_per_s(ks, frames=1, hw=(H_LR, W_LR)):
    """The CPU restatement of the reference's generator (oracle/) on all host cores; bounded sample."""
    import numpy as np
    import torch
    from oracle import srgan_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    w = O.to_torch(O.init_upscaler_orig_weights(seed=1234, kernel_size=ks))
    rng = np.random.default_rng(0)
    with torch.no_grad():
        O.upscaler_orig_forward(w, rng.uniform(-1, 1, (1, 32, 32, 3)).astype(np.float32))  # warm oneDNN
        x = rng.uniform(-1, 1, (frames, hw[0], hw[1], 3)).astype(np.float32)
        t0 = time.perf_counter()
        O.upscaler_orig_forward(w, x)
        dt = time.perf_counter() - t0
    return frames / dt, cores, dt


def run_reference(args, rank, world):
    """Reference arm: the reference's own (CPU) implementation of the path on the host cores. TensorFlow/Keras are
    not installable here, so this times the oracle restatement (kind = 'port')."""
    if rank != 0:
        return
    vals = []
    cores = os.cpu_count() or 1
    for i in range(args.warmup + args.steps):
        fps, cores, dt = cpu_oracle_frames_per_s(args.kernel_size, 1)
        if i >= args.warmup:
            vals.append(dt)
        if sum(vals) > 240:  # keep the whole run within a few minutes
            break
    steps = len(vals)
    ms = 1000.0 * sum(vals) / max(steps, 1)
    v = 1000.0 / ms
    line = {"impl": "reference", "metric": "x4 frames/sec 270p->1080p", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "generator x4 inference 270x480->1080x1920, k=%d, CPU oracle, 1 frame/step" % args.kernel_size},
            "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
                             "sample": "1 frame of 270x480 per step (PyTorch-CPU fp32 restatement; reference TensorFlow unavailable on this box)"},
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_oracle_train_samples_per_s(ks, B=2, hw=(96, 96)):
    """the CPU oracle's generator-only training step (fp32 autograd + Adam) on all host cores, at a
    reduced batch (B = 2 patches; a batch of 16 needs ~15 GB and > 15 minutes of CPU backward); samples/s."""
    import numpy as np
    import torch
    from oracle import srgan_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    h, w_ = hw
    ref = O.GeneratorTrainer(O.init_upscaler_orig_weights(seed=1234, kernel_size=ks), O.init_vgg19_weights(), "vgg-mse", 1e-3)
    rng = np.random.default_rng(0)
    hr = rng.uniform(-1, 1, (B, 4 * h, 4 * w_, 3)).astype(np.float32)
    x = hr.reshape(B, h, 4, w_, 4, 3).mean(axis=(2, 4)).astype(np.float32)
    t0 = time.perf_counter()
    ref.train_on_batch(x, hr)
    dt = time.perf_counter() - t0
    return {"value": B / dt, "unit": "samples/s", "cores": cores, "kind": "port", "like_for_like_batch": False,
            "sample": "one train_on_batch of %d patches of %dx%d -> %dx%d (%.1f s) through the PyTorch-CPU fp32 autograd oracle "
                      "(the step named by the config has 16: scaled by batch size, BatchNorm over 2 instead of 16 samples)" % (B, h, w_, 4 * h, 4 * w_, dt)}


def cpu_oracle_gan_images_per_s(ks, n=1, lr=128):
    """one train_gan.py loop iteration (predict + discriminator step + GAN step) of the CPU oracle, n LR
    images of 128x128 -> 512x512."""
    import numpy as np
    import torch
    from oracle import srgan_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ref = O.GanTrainer(O.init_upscaler_orig_weights(seed=1234, kernel_size=ks),
                       O.init_discriminator_simple_512_weights(seed=77, input_shape=(4 * lr, 4 * lr, 3)), O.init_vgg19_weights(), "vgg-mse", 1e-3, 1.0, 1e-5)
    rng = np.random.default_rng(0)
    hr = rng.uniform(-1, 1, (n, 4 * lr, 4 * lr, 3)).astype(np.float32)
    x = hr.reshape(n, lr, 4, lr, 4, 3).mean(axis=(2, 4)).astype(np.float32)
    t0 = time.perf_counter()
    ref.step(x, hr)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": "samples/s", "cores": cores, "kind": "port", "like_for_like_batch": False,
            "sample": "one loop iteration with n = %d LR image(s) of %dx%d -> %dx%d (%.1f s) through the PyTorch-CPU fp32 oracle "
                      "(the GPU leg pushes 6 per iteration)" % (n, lr, lr, 4 * lr, 4 * lr, dt)}


def training_path_text(compiled_list, dev, world):
    """Words for a training leg's `config`, derived from the path the timed steps actually took: CUDA-graph replay or eager
    launches, and for N > 1 how the BatchNorm sums and the gradients crossed the replicas."""
    from upscaler import training as T
    replay = all(bool(c._graphs) and not c._graph_off for c in compiled_list)
    txt = ", steps replayed from captured CUDA graphs" if replay else ", eager launches (no CUDA-graph replay)"
    if world > 1:
        sync = T._make_sync()
        peer = sync.peer_for(dev) if sync is not None else None
        bn = ("per-replica BatchNorm statistics" if sync is None else
              "BatchNorm sums exchanged over NVLink peer memory inside the finalize kernels" if peer is not None else
              "BatchNorm sums all-reduced over NCCL")
        gr = ("one NCCL all-reduce of the flat gradient after backward (not overlapped)" if replay else
              "bucketed NCCL gradient all-reduce overlapped with backward")
        txt += ", data-parallel: %s; %s" % (bn, gr)
    return txt


def bench_train_step(args, rank, world, dev, barrier, B=16, hw=(96, 96), builder="orig", loss_kind="vgg-mse", calls=1,
                     what="generator-only step"):
    """Generator-only training step with the VGG19 content loss (train.py; BASELINE.json configs[2]): B LR patches of
    hw per GPU through compile_training_model(...).train_on_batch(host x, host y) -- forward with batch-statistics
    BatchNorm, VGG19 loss, backward, Adam; data-parallel over NCCL when N > 1 (cross-replica BatchNorm sums, bucketed
    gradient all-reduce overlapped with backward).  H2D of the batch and the D2H read of the loss are inside.
    `calls` > 1 = one ITERATION of train_cgc.py:275-277: that many train_on_batch calls on different LR renderings of the
    same HR frame (gen1, gen2, scaled), each a full step; value then counts HR frames/s."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from upscaler import ops
    from upscaler import model as M
    ks = args.kernel_size
    h, w_ = hw
    make = {"orig": M.make_upscaler_orig, "skip-con": M.make_upscaler_skip_con}[builder]
    up = make((4 * h, 4 * w_, 3), kernel_size=ks)
    lobj = {"vgg-mse": lambda: M.VGG_MSE_LOSS((4 * h, 4 * w_, 3), 1e-3), "vgg": lambda: M.VGG_LOSS((4 * h, 4 * w_, 3))}[loss_kind]()
    tm = M.compile_training_model(up, lobj.loss)
    rng = np.random.default_rng(100 + rank)
    hr = rng.uniform(-1, 1, (B, 4 * h, 4 * w_, 3)).astype(np.float32)
    base = hr.reshape(B, h, 4, w_, 4, 3).mean(axis=(2, 4)).astype(np.float32)
    xs = [np.clip(base + 0.05 * c * rng.standard_normal(base.shape).astype(np.float32), -1, 1) for c in range(calls)]
    for _ in range(6):      # warm-up past training.GRAPH_WARMUP: the one-off CUDA-graph capture of the step is not timed
        for x in xs:
            loss = tm.train_on_batch(x, hr)
    barrier()
    ops.reset_launch_count()
    from upscaler import training as _tr
    _tr.HOST_STATS.update(wait_s=0.0, reads=0, stage_wait_s=0.0)
    t0 = time.perf_counter()
    for _ in range(args.train_steps):
        for x in xs:
            loss = tm.train_on_batch(x, hr)       # returns the loss as a host float: the step's D2H read + sync
    host_wait_ms = 1000.0 * _tr.HOST_STATS["wait_s"] / args.train_steps
    stage_wait_ms = 1000.0 * _tr.HOST_STATS.get("stage_wait_s", 0.0) / args.train_steps
    barrier()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = 1000.0 * float(t.item()) / args.train_steps
    peaks = load_peaks()
    # 3 G + 3 V per sample; V scales with the HR pixel count (114.66 GFLOP at 384 x 384)
    flops = B * 3 * (GEN_FLOPS[ks] * (h * w_) / float(H_LR * W_LR) + 114.66e9 * (16 * h * w_) / float(384 * 384))
    if (B, hw, builder) == (16, (96, 96), "orig"):
        flops = TRAIN_STEP_FLOPS[ks]
    flops *= calls       # (skip-con: the generator FLOPs are make_upscaler_orig's; its 224-filter 3x3 ConvT layers do less -> a lower bound on the fraction is not claimed)
    tf = flops * world / (ms * 1e-3) / 1e12
    out = {"metric": "train samples/s", "value": world * B * 1000.0 / ms, "unit": "samples/s" if calls == 1 else "HR frames/s (iterations/s x batch)",
           "ms_per_step": ms, "steps": args.train_steps,
           "config": "%s, %s content loss, %d x (%dx%d -> %dx%d) per GPU, make_upscaler_%s k=%d, "
           "%s storage / fp32 accumulate / fp32 master weights + Adam%s" % (what, loss_kind, B, h, w_, 4 * h, 4 * w_, builder.replace("-", "_"), ks, str(up.net.train_dtype).replace("torch.", ""), training_path_text([tm._compiled], dev, world)),
           "train_on_batch_calls_per_step": calls,
           "host_wait_ms_per_step": host_wait_ms,
           "host_stage_wait_ms_per_step": stage_wait_ms,      # time the host waited for its staging threads (pageable -> pinned copy of the batch)      # time rank 0's host sat in the loss read: ~0 = bound by the host's launch rate
           "gpu_launches_per_step": ops.launch_count() // args.train_steps, "loss": float(loss),
           "h2d_bytes_per_step": int(sum(x.nbytes for x in xs) + calls * hr.nbytes), "d2h_bytes_per_step": 4 * calls}
    if builder == "orig":
        out.update({"algorithmic_tflops": tf, "frac_of_bf16_sustained_peak": tf / world / peaks["bf16_tflops_sustained"]})
    del tm, up
    torch.cuda.empty_cache()
    return out


GAN_STEP_FLOPS_PER_IMAGE = {3: 4 * 155.1e9 + 8 * 36.34e9 + 3 * 203.84e9, 5: 2.39e12}   # 4G + 8D + 3V at HR 512^2, SURVEY.md 8(d)


def bench_gan_step(args, rank, world, dev, barrier):
    """Full SRGAN adversarial step (BASELINE.json configs[3]; train_gan.py:296-321 with its defaults HR 512x512, bs 2 -> n = 6
    LR images of 128x128): gen_train.predict(x) -> host concat [real, fake] -> disc_train.train_on_batch -> gan_train
    .train_on_batch(x, [hr, +1]) through make_and_compile_gan(...), VGG19 content loss + Wasserstein adversarial loss, one
    shared Adam; host arrays in and out exactly as the reference's loop moves them; NCCL data-parallel when N > 1."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from upscaler import ops
    from upscaler.model import (VGG_MSE_LOSS, make_and_compile_gan, make_discriminator_simple_512, make_upscaler_orig,
                                wasserstein_loss)
    n, lr, ks = 6, 128, args.kernel_size
    hr_shape = (4 * lr, 4 * lr, 3)
    G = make_upscaler_orig(hr_shape, kernel_size=ks)
    D = make_discriminator_simple_512(hr_shape, activation="none")
    gen_train, disc_train, gan_train = make_and_compile_gan(G, D, (lr, lr, 3), hr_shape, VGG_MSE_LOSS(hr_shape, 1e-3).loss, 1,
                                                           wasserstein_loss, 1e-5)
    rng = np.random.default_rng(200 + rank)
    hr = rng.uniform(-1, 1, (n,) + hr_shape).astype(np.float32)
    x = hr.reshape(n, lr, 4, lr, 4, 3).mean(axis=(2, 4)).astype(np.float32)
    d_y = np.concatenate((np.ones(n), -np.ones(n))).astype(np.float32)
    ones = np.ones(n, np.float32)

    def step():
        gen = gen_train.predict(x)
        ld = disc_train.train_on_batch(np.concatenate((hr, gen), axis=0), d_y)
        lg = gan_train.train_on_batch(x, [hr, ones])
        return ld, lg

    for _ in range(6):      # past training.GRAPH_WARMUP: the one-off CUDA-graph captures are not timed
        step()
    barrier()
    ops.reset_launch_count()
    steps = max(3, args.train_steps // 2)
    t0 = time.perf_counter()
    for _ in range(steps):
        ld, lg = step()
    barrier()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = 1000.0 * float(t.item()) / steps
    peaks = load_peaks()
    tf = GAN_STEP_FLOPS_PER_IMAGE[ks] * n * world / (ms * 1e-3) / 1e12
    out = {"metric": "GAN train LR images/s", "value": world * n * 1000.0 / ms, "unit": "samples/s", "ms_per_step": ms, "steps": steps,
           "config": "train_gan.py step: predict + discriminator step (2n images) + GAN step, n = %d LR images of 128x128 -> 512x512 per GPU, "
                     "k=%d, simple_512 discriminator, vgg-mse + Wasserstein, shared Adam, %s training storage%s" % (n, ks, str(G.net.train_dtype).replace("torch.", ""), training_path_text([disc_train._compiled, gan_train._compiled], dev, world)),
           "algorithmic_tflops": tf, "frac_of_bf16_sustained_peak": tf / world / peaks["bf16_tflops_sustained"],
           "gpu_launches_per_step": ops.launch_count() // steps, "loss_disc": float(ld), "loss_gan": [float(v) for v in lg]}
    del gen_train, disc_train, gan_train, G, D
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=8, help="frames per step per GPU")
    ap.add_argument("--kernel-size", type=int, default=3, dest="kernel_size")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the training-step measurement (config 3)")
    ap.add_argument("--train-steps", type=int, default=8, dest="train_steps")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback for the product path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        if os.environ.get("OMP_NUM_THREADS") == "1" and not os.environ.get("B2_KEEP_OMP"):
            # torchrun's default of one host thread per rank makes torch's pageable -> pinned staging copies single-threaded
            # (30 MB: 4.8 ms instead of 1.1 ms); give every rank its share of the host cores instead
            torch.set_num_threads(max(1, min(16, (os.cpu_count() or 1) // world)))

    from upscaler import ops
    from upscaler.model import make_upscaler_orig

    B, ks = args.batch, args.kernel_size
    model = make_upscaler_orig((4 * H_LR, 4 * W_LR, 3), kernel_size=ks)
    net = model.net
    gen = torch.Generator(device="cpu").manual_seed(1000 + rank)   # each rank upscales its own shard of frames
    x_host = (torch.rand((B, H_LR, W_LR, 3), generator=gen) * 2 - 1).numpy()
    x_dev = torch.from_numpy(x_host).to(dev)
    out_dev = torch.empty((B, 4 * H_LR, 4 * W_LR, 3), dtype=torch.float32, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ------------------------------------------------------------------ device-resident throughput
    for _ in range(args.warmup):
        net.forward_infer(x_dev, out_dev)
    barrier()
    ops.reset_launch_count()
    net.layer_events = {}            # per-layer CUDA events on the launching stream (roofline)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        net.forward_infer(x_dev, out_dev)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    launches = ops.launch_count()
    layer_ms = {k: sum(a.elapsed_time(b) for a, b in v) / len(v) for k, v in net.layer_events.items()}
    net.layer_events = None
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = world * B * 1000.0 / ms_step

    # ------------------------------------------------------------------ end to end through Model.predict (host buffers)
    y = None
    for _ in range(3):      # holds the previous result like the timed loop does, so both pinned result blocks exist
        y = model.predict(x_host, batch_size=B)
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(e2e_steps):
        y = model.predict(x_host, batch_size=B)
    barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_value = world * B * e2e_steps / float(t_e2e.item())
    assert y.shape == (B, 4 * H_LR, 4 * W_LR, 3)
    # same pipeline with uint8 frames at both ends (the reference's frame-upscaling loop, data.py:358-363): the
    # uint8 <-> float conversions run on the device, PCIe carries a quarter of the bytes
    f_host = np.random.default_rng(rank).integers(0, 256, (B, H_LR, W_LR, 3), dtype=np.uint8)
    yu = None
    for _ in range(3):
        yu = model.predict_uint8(f_host, batch_size=B)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        yu = model.predict_uint8(f_host, batch_size=B)
    barrier()
    t_u8 = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_u8, op=dist.ReduceOp.MAX)
    e2e_u8 = world * B * e2e_steps / float(t_u8.item())

    # ------------------------------------------------------------------ training step (BASELINE.json configs[2])
    train = train_gan = train_full = train_cgc = train_cgc_mini = train_strong = None
    if not args.no_train:
        train = bench_train_step(args, rank, world, dev, barrier)
        if world > 1 and 16 % world == 0:
            # strong scaling of the same step: the config's global batch of 16 split over the ranks
            train_strong = bench_train_step(args, rank, world, dev, barrier, B=16 // world,
                                            what="generator-only step, STRONG scaling: global batch 16 = %d per GPU" % (16 // world))
        train_gan = bench_gan_step(args, rank, world, dev, barrier)
        # train.py's own workload: batch 1 of a full 270x480 -> 1080x1920 frame
        train_full = bench_train_step(args, rank, world, dev, barrier, B=1, hw=(H_LR, W_LR))
        # BASELINE.json configs[4]: one train_cgc.py ITERATION (train_cgc.py:275-277) = three full-frame train_on_batch calls
        # (gen1 / gen2 / scaled renderings -> the same HR frame), make_upscaler_skip_con, its default loss vgg-only, bs 1
        train_cgc = bench_train_step(args, rank, world, dev, barrier, B=1, hw=(H_LR, W_LR), builder="skip-con", loss_kind="vgg",
                                     calls=3, what="train_cgc.py iteration (3 train_on_batch calls)")
        # train_cgc_mini.py:269-275: the three renderings concatenated into ONE batch of 3 x bs (bs = 3) 128x128 -> 512x512 crops
        train_cgc_mini = bench_train_step(args, rank, world, dev, barrier, B=9, hw=(128, 128), builder="skip-con", loss_kind="vgg",
                                          what="train_cgc_mini.py step (3 x bs = 9 crops in one batch)")
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            train["cpu_baseline"] = cpu_oracle_train_samples_per_s(ks)
            train_gan["cpu_baseline"] = cpu_oracle_gan_images_per_s(ks)

    if rank == 0:
        peaks = load_peaks()
        dom = "upscaling/1/block/conv_transp"
        dom_ms = layer_ms.get(dom)
        achieved = CONVT2_FLOPS[ks] * B / (dom_ms * 1e-3) / 1e12 if dom_ms else None
        # DRAM bytes of that launch from the committed `ncu --set full` summary (captured at B = 8, k = 3; algorithmic =
        # 2.12 GB input + 8.49 GB output per 8 frames)
        prof = profiled_dram_traffic() if ks == 3 else None
        traffic = prof[0] * B / 8 if prof else None
        roof = {"bound": "tensor", "kernel": "conv_phases2_kernel (all 4 sub-pixel phases of " + dom + " in one persistent CTA-pair launch)",
                "achieved": achieved, "peak": peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
                "frac": (achieved / peaks["bf16_tflops_sustained"]) if achieved else None,
                "peak_kind": "bf16 dense sustained, %s" % peaks["source"], "traffic": traffic,
                "traffic_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)",
                "traffic_source": ("profiles/" + prof[1]) if prof else None,
                "traffic_profiled_offline": True,    # read from the committed ncu capture (B = 8) and rescaled by batch; NOT measured in this run
                "algorithmic_bytes": (2.12e9 + 8.49e9) * B / 8 if ks == 3 else None,
                "share_of_step": (dom_ms / ms_step) if dom_ms else None}
        whole = GEN_FLOPS[ks] * world * B / (ms_step * 1e-3) / 1e12
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            fps, cores, dt = cpu_oracle_frames_per_s(ks, 1)
            cpu = {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                   "sample": "1 frame of 270x480 (%.1f s) through the PyTorch-CPU fp32 oracle; reference TensorFlow unavailable on this box" % dt}
        line = {"metric": "x4 frames/sec 270p->1080p", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "fp16 storage / fp32 accumulate (tcgen05 kind::f16)", "data": "synthetic",
                "config": {"workload": "SRGAN generator x4 inference, %d frames/step/GPU of 270x480 -> 1080x1920, k=%d, frames sharded, no collective" % (B, ks),
                           "l2": "working set per step (%.1f GB of activations) >> 126 MB L2" % (B * 1.4),
                           "whole_step_tflops": whole, "frac_of_bf16_sustained_peak": whole / world / peaks["bf16_tflops_sustained"]},
                "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": int(x_host.nbytes),
                        "d2h_bytes_per_step": int(y.nbytes), "api": "upscaler Model.predict(ndarray)->ndarray"},
                "e2e_uint8": {"value": e2e_u8, "unit": "frames/s", "h2d_bytes_per_step": int(f_host.nbytes),
                              "d2h_bytes_per_step": int(yu.nbytes),
                              "api": "Model.predict_uint8(uint8 frames)->uint8 frames (data.py:253-270 conversions fused on the device)"},
                "gpu_launches": launches,
                "host": {"cpu_count": os.cpu_count(), "affinity": len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else None,
                         "torch_threads": torch.get_num_threads()},
                "roofline": roof, "cpu_baseline": cpu, "train": train, "train_strong_scaling": train_strong, "train_gan": train_gan, "train_full_frame": train_full,
                "train_cgc_iteration": train_cgc, "train_cgc_mini": train_cgc_mini,
                "layers_ms": {k: round(v, 4) for k, v in sorted(layer_ms.items(), key=lambda kv: -kv[1])[:8]}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
