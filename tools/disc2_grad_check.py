"""Diagnostic: is the two-pass discriminator gradient (train_gan3.py:353, step_disc2) wrong, or is its direction lost in the
cancellation  g(real) - g(fake)  under 16-bit storage?  For bf16 and fp16 critic storage: each pass's gradient alone
(coefficients (1, 0) and (0, -1)) and the sum, per tensor against fp32 autograd."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np
import torch

from oracle import srgan_oracle as O
from upscaler import ops
from upscaler.networks import DiscriminatorSimple512Net

HR, n = 64, 8
rng = np.random.default_rng(16)
real = rng.uniform(-1, 1, (n, HR, HR, 3)).astype(np.float32)
real = ((real + np.roll(real, 1, 1) + np.roll(real, 1, 2)) / 3).astype(np.float32)
fake = np.clip(real + 0.2 * rng.standard_normal(real.shape).astype(np.float32), -1, 1)
dw = O.init_discriminator_simple_512_weights(seed=3, input_shape=(HR, HR, 3))


def oracle(cr, cf):
    w = O.to_torch(dw, requires_grad=True)
    r = O.discriminator_simple_512_forward(w, real, "none", True, [])
    f = O.discriminator_simple_512_forward(w, fake, "none", True, [])
    (cr * r.mean() + cf * f.mean()).backward()
    return {k: [None if t.grad is None else t.grad.numpy().copy() for t in ts] for k, ts in w.items()}


def ours(dtype, cr, cf):
    net = DiscriminatorSimple512Net((HR, HR, 3), weights=dw, dtype=dtype)
    rd, fd = torch.from_numpy(real).cuda(), torch.from_numpy(fake).cuda()
    sr, tr = net.forward(rd, True)
    sf, tf = net.forward(fd, True)
    net.ensure_grads().zero_()
    S = 256.0 * n if dtype == torch.float16 else 1.0
    for scores, tape, coef in ((sr, tr, cr), (sf, tf, cf)):
        if coef == 0:
            continue
        ds = torch.empty_like(scores)
        ops.mean_prod(scores, None, None, 0.0, coef * S, ds, 1.0)
        net.backward(tape, ds, True, False, None)
    torch.cuda.synchronize()
    return net, {k: [g.cpu().numpy() / S for g in gs] for k, gs in net.grads.items()}


for dtype in (torch.bfloat16, torch.float16):
    for cr, cf, tag in ((1.0, 0.0, "real pass alone"), (0.0, -1.0, "fake pass alone"), (1.0, -1.0, "both (the step)")):
        ref = oracle(cr, cf)
        net, got = ours(dtype, cr, cf)
        rows = []
        for name, kind, cfg in net.specs:
            for j, g in enumerate(got[name]):
                r = ref[name][j]
                if r is None or (kind in ("conv", "dense") and j == 1 and name != "discriminator/final/Dense_3"):
                    continue
                g = g.reshape(r.shape)
                if np.linalg.norm(r) < 1e-9:
                    continue
                rows.append((name, j, float(np.linalg.norm(g) / np.linalg.norm(r)),
                             float(np.vdot(g, r) / max(np.linalg.norm(g) * np.linalg.norm(r), 1e-30)), float(np.linalg.norm(r))))
        worst = min(rows, key=lambda t: t[3])
        print("%s  %-16s worst cos %.4f (%s[%d], norm ratio %.3f, |ref| %.3e); median cos %.4f" % (
            str(dtype).replace("torch.", ""), tag, worst[3], worst[0], worst[1], worst[2], worst[4], float(np.median([t[3] for t in rows]))))
        if tag.startswith("both"):
            for t in rows[:6]:
                print("      %-40s[%d] norm ratio %.3f cos %.4f |ref| %.3e" % t)
