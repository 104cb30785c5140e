"""Development aid: times the generator-only training step (BASELINE config 3: B=16, LR 96x96 -> HR 384x384, vgg-mse)
through the public API (compile_training_model + train_on_batch)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np, torch
from upscaler.model import VGG_MSE_LOSS, compile_training_model, make_upscaler_orig
from upscaler import ops
if os.environ.get('B2_NO_GRAPH'):
    from upscaler import training
    training.GRAPH_STEPS = False      # A/B: eager launches instead of CUDA-graph replay
if os.environ.get('B2_NO_PDL'):
    ops.set_pdl(False)      # A/B: plain stream order instead of programmatic dependent launch
ks = int(sys.argv[1]) if len(sys.argv) > 1 else 3
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
lr_hw = int(sys.argv[4]) if len(sys.argv) > 4 else 96
up = make_upscaler_orig((4 * lr_hw, 4 * lr_hw, 3), kernel_size=ks)
tm = compile_training_model(up, VGG_MSE_LOSS((4 * lr_hw, 4 * lr_hw, 3), 1e-3).loss)
rng = np.random.default_rng(0)
hr = rng.uniform(-1, 1, (B, 4 * lr_hw, 4 * lr_hw, 3)).astype(np.float32)
x = hr.reshape(B, lr_hw, 4, lr_hw, 4, 3).mean(axis=(2, 4)).astype(np.float32)
for _ in range(6):      # past training.GRAPH_WARMUP: the one-off CUDA-graph capture stays out of the timed steps
    l = tm.train_on_batch(x, hr)
torch.cuda.synchronize()
ops.reset_launch_count()
t0 = time.perf_counter()
for _ in range(steps):
    l = tm.train_on_batch(x, hr)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / steps
fl = {3: 9.69e12, 5: 15.55e12}[ks] * B / 16 * (lr_hw / 96.0) ** 2
print("k=%d B=%d LR %d: %.2f ms/step, %.1f samples/s, %.1f TFLOP/s algorithmic, loss %.5f, %d launches/step, peak mem %.1f GB" % (
    ks, B, lr_hw, dt * 1e3, B / dt, fl / dt / 1e12, l, ops.launch_count() // steps, torch.cuda.max_memory_allocated() / 2 ** 30))
