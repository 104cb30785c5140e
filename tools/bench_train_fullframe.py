"""Development aid: the generator-only training step on ONE full frame (train.py's real workload, SURVEY.md D6 / configs[4]:
x (1,270,480,3) -> y (1,1080,1920,3), vgg-mse content loss, Adam) through compile_training_model + train_on_batch."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np, torch
from upscaler.model import VGG_MSE_LOSS, compile_training_model, make_upscaler_orig
from upscaler import ops
ks = int(sys.argv[1]) if len(sys.argv) > 1 else 3
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
H, W = 270, 480
up = make_upscaler_orig((4 * H, 4 * W, 3), kernel_size=ks)
tm = compile_training_model(up, VGG_MSE_LOSS((4 * H, 4 * W, 3), 1e-3).loss)
rng = np.random.default_rng(0)
hr = rng.uniform(-1, 1, (1, 4 * H, 4 * W, 3)).astype(np.float32)
x = hr.reshape(1, H, 4, W, 4, 3).mean(axis=(2, 4)).astype(np.float32)
for _ in range(2):
    l = tm.train_on_batch(x, hr)
torch.cuda.synchronize()
ops.reset_launch_count()
t0 = time.perf_counter()
for _ in range(steps):
    l = tm.train_on_batch(x, hr)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / steps
fl = 3 * {3: 1227.09e9, 5: 2942.77e9}[ks] + 3 * 1611.3e9
print("full-frame train step k=%d: %.2f ms/step, %.2f frames/s, %.1f TFLOP/s algorithmic, loss %.5f, %d launches, peak mem %.1f GB" % (
    ks, dt * 1e3, 1 / dt, fl / dt / 1e12, l, ops.launch_count() // steps, torch.cuda.max_memory_allocated() / 2 ** 30))
