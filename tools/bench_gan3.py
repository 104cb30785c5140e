"""Development aid: times one train_gan3.py loop iteration (train_gan3.py:339-354: predict -> disc_train.train_on_batch([hr, gen])
-> gan_train.train_on_batch([x, hr], [hr, -1])) through make_and_compile_gan2 with RelativisticLosses('log-sigm'), n = 6 LR
images of 128x128 -> 512x512.  B2_NO_GRAPH=1: eager launches instead of CUDA-graph replay."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np, torch
from upscaler import ops, training
from upscaler.model import (RelativisticLosses, VGG_MSE_LOSS, make_and_compile_gan2, make_discriminator_simple_512,
                            make_upscaler_orig)
if os.environ.get("B2_NO_GRAPH"):
    training.GRAPH_STEPS = False
n, lr = 6, 128
hr_shape = (4 * lr, 4 * lr, 3)
G = make_upscaler_orig(hr_shape, kernel_size=3)
D = make_discriminator_simple_512(hr_shape, activation="none")
gen_train, disc_train, gan_train = make_and_compile_gan2(G, D, (lr, lr, 3), hr_shape, VGG_MSE_LOSS(hr_shape, 1e-3).loss, 1,
                                                        lambda: RelativisticLosses(loss_activation="log-sigm"), 1e-5)
rng = np.random.default_rng(0)
hr = rng.uniform(-1, 1, (n,) + hr_shape).astype(np.float32)
x = hr.reshape(n, lr, 4, lr, 4, 3).mean(axis=(2, 4)).astype(np.float32)
fake_y = -np.ones(n, np.float32)


def step(rec=None):
    t = [time.perf_counter()]
    gen = gen_train.predict(x); t.append(time.perf_counter())
    ld = disc_train.train_on_batch([hr, gen], fake_y); t.append(time.perf_counter())
    lg = gan_train.train_on_batch([x, hr], [hr, fake_y]); t.append(time.perf_counter())
    if rec is not None:
        rec.append([1e3 * (b - a) for a, b in zip(t, t[1:])])
    return ld, lg


for _ in range(6):
    step()
rec = []
ops.reset_launch_count()
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
for _ in range(steps):
    ld, lg = step(rec)
r = np.array(rec).mean(0)
print("train_gan3.py iteration (%s): predict %.2f ms, disc step %.2f, gan step %.2f; total %.2f ms; %d launches/step; losses %.4f %s" % (
    "eager" if not training.GRAPH_STEPS else "CUDA-graph replay", r[0], r[1], r[2], r.sum(), ops.launch_count() // steps, ld,
    np.round(lg, 4)))
