"""Development aid: times the adversarial step of bench.py's train_gan (configs[3]) piece by piece on the host side."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np, torch
from upscaler import ops
if os.environ.get('B2_NO_GRAPH'):
    training_mod = __import__('upscaler.training', fromlist=['x'])
    training_mod.GRAPH_STEPS = False      # A/B: eager launches instead of CUDA-graph replay
if os.environ.get('B2_NO_PDL'):
    ops.set_pdl(False)      # A/B: plain stream order instead of programmatic dependent launch
from upscaler import model as M
from upscaler.model import VGG_MSE_LOSS, make_and_compile_gan, make_upscaler_orig, wasserstein_loss
ks = int(sys.argv[1]) if len(sys.argv) > 1 else 3
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
disc = sys.argv[3] if len(sys.argv) > 3 else "simple"          # simple | thin | sparse
n, lr = 6, 128
hr_shape = (4 * lr, 4 * lr, 3)
G = make_upscaler_orig(hr_shape, kernel_size=ks)
D = getattr(M, "make_discriminator_%s_512" % disc)(hr_shape, activation="none")
gen_train, disc_train, gan_train = make_and_compile_gan(G, D, (lr, lr, 3), hr_shape, VGG_MSE_LOSS(hr_shape, 1e-3).loss, 1, wasserstein_loss, 1e-5)
rng = np.random.default_rng(0)
hr = rng.uniform(-1, 1, (n,) + hr_shape).astype(np.float32)
x = hr.reshape(n, lr, 4, lr, 4, 3).mean(axis=(2, 4)).astype(np.float32)
d_y = np.concatenate((np.ones(n), -np.ones(n))).astype(np.float32)
ones = np.ones(n, np.float32)
def step(rec=None):
    t = [time.perf_counter()]
    gen = gen_train.predict(x); t.append(time.perf_counter())
    d_in = np.concatenate((hr, gen), axis=0); t.append(time.perf_counter())
    ld = disc_train.train_on_batch(d_in, d_y); t.append(time.perf_counter())
    lg = gan_train.train_on_batch(x, [hr, ones]); t.append(time.perf_counter())
    if rec is not None:
        rec.append([1e3 * (b - a) for a, b in zip(t, t[1:])])
from upscaler import training
_up = [0.0]
_orig_finish, _orig_begin = training._uploader.finish, training._uploader.begin
def _begin(*a, **k):
    t0 = time.perf_counter(); r = _orig_begin(*a, **k); _up[0] += time.perf_counter() - t0; return r
def _finish(*a, **k):
    t0 = time.perf_counter(); r = _orig_finish(*a, **k); _up[0] += time.perf_counter() - t0; return r
training._uploader.begin, training._uploader.finish = _begin, _finish
for _ in range(6):      # past training.GRAPH_WARMUP
    step()
_up[0] = 0.0
rec = []
ops.reset_launch_count()
for _ in range(steps):
    step(rec)
r = np.array(rec).mean(0)
print("host staging (begin+finish) per step: %.2f ms" % (1e3 * _up[0] / steps))
print("GAN step k=%d: predict %.2f ms, host concat %.2f, disc step %.2f, gan step %.2f; total %.2f ms; %d launches/step" % (
    ks, r[0], r[1], r[2], r[3], r.sum(), ops.launch_count() // steps))
