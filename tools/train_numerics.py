"""Where does the gradient noise of the training step come from?  (VERDICT r1 "Next round" 1c.)  Single GPU.

For each storage configuration of the generator's training passes -- bf16 (S = 1) and fp16 with gradient scaling
(training.GRAD_SCALE) -- on the dp_check problem (2 residual blocks, 8 patches 32x32 -> 128x128, vgg-mse) and on the full
16-block generator (4 patches 32x32):
  * run-to-run: the same step twice on the same data (what is left once the BatchNorm sums are deterministic: the
    red.global.add flushes of the split-K weight-gradient kernels);
  * reversed batch: the same step with the samples in reverse order (identical mathematics; decorrelates every 16-bit
    rounding because the BatchNorm sums change in the last bit) = the storage-rounding floor of each gradient tensor;
  * against the fp32 autograd oracle: per-tensor relative L2 error / cosine of the first step's gradients, and the loss
    of 4 free-running steps.
Prints one table per configuration; the numbers quoted in DESIGN.md 6 come from profiles/r02_*_train_numerics.log."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np
import torch

from oracle import srgan_oracle as O


def data(B, lr, step):
    rng = np.random.default_rng(500 + step)
    hr = rng.uniform(-1, 1, (B, 4 * lr, 4 * lr, 3)).astype(np.float32)
    hr = ((hr + np.roll(hr, 1, 1) + np.roll(hr, 1, 2) + np.roll(hr, (1, 1), (1, 2))) / 4).astype(np.float32)
    return hr.reshape(B, lr, 4, lr, 4, 3).mean(axis=(2, 4)).astype(np.float32), hr


def build(rbn, lr, dtype, gw, vw):
    from upscaler import networks
    from upscaler.keras_compat import Adam, Model
    from upscaler.model import VGG_MSE_LOSS, compile_training_model
    net = networks.UpscalerOrigNet(kernel_size=3, res_block_num=rbn, train_dtype=dtype)
    up = Model(net=net, input_shape=(lr, lr, 3), name="model_1")
    up.net.set_weights(gw)
    tm = compile_training_model(up, VGG_MSE_LOSS((4 * lr, 4 * lr, 3), 1e-3, vgg19=vw).loss, optimizer=Adam())
    return up, tm


def first_grad(up, tm, x, hr):
    loss = tm.train_on_batch(x, hr)
    torch.cuda.synchronize()
    S = tm._compiled._scale_for(up.net, hr.size, hr.shape[0])
    return loss, up.net.grad_flat.detach().cpu().numpy().astype(np.float64) / S


def per_tensor(net, a, b):
    out = {}
    for (name, j), (o, n, shp) in net.offsets.items():
        if o >= net.n_trainable:
            continue
        x, y = a[o:o + n], b[o:o + n]
        ny = float(np.linalg.norm(y))
        if ny < 1e-6 * np.sqrt(n) * max(1e-30, float(np.abs(b).max())) or ny == 0:
            continue
        kind = dict((nm, k) for nm, k, _ in net.specs)[name]
        if kind == "conv" and j == 1 and name not in ("initial/conv", "final/conv"):
            continue                      # bias in front of a train-mode BatchNorm: analytically zero
        out["%s[%d]" % (name, j)] = (float(np.linalg.norm(x - y)) / ny,
                                     float(np.vdot(x, y) / max(float(np.linalg.norm(x)) * ny, 1e-300)))
    return out


def summarize(tag, d):
    rel = np.array([v[0] for v in d.values()])
    cos = np.array([v[1] for v in d.values()])
    worst = max(d.items(), key=lambda kv: kv[1][0])
    print("    %-28s rel L2: median %.4f  p90 %.4f  worst %.4f (%s);  worst cos %.5f" % (
        tag, np.median(rel), np.percentile(rel, 90), rel.max(), worst[0], cos.min()))
    return rel.max()


def main():
    torch.cuda.set_device(0)
    from upscaler import training
    vw = O.init_vgg19_weights()
    for rbn, B, lr in ((2, 8, 32), (16, 4, 32)):
        gw = O.init_upscaler_orig_weights(seed=11, kernel_size=3, res_block_num=rbn)
        x, hr = data(B, lr, 0)
        ref = O.GeneratorTrainer(gw, vw, "vgg-mse", 1e-3, res_block_num=rbn)
        w = O.to_torch(gw, requires_grad=True)
        y = O.upscaler_orig_forward(w, x, True, [], 4, rbn)
        O.content_loss("vgg-mse", O.to_torch(vw), torch.from_numpy(hr), y, 1e-3).backward()
        ref_losses = [ref.train_on_batch(*data(B, lr, s)) for s in range(4)]
        for dtype, gs in ((torch.bfloat16, None), (torch.float16, "numel"), (torch.float16, 65536.0)):
            if gs is not None:
                training.GRAD_SCALE["G"] = gs
            up, tm = build(rbn, lr, dtype, gw, vw)
            net = up.net
            gref = np.zeros(net.n_trainable)
            for (name, j), (o, n, shp) in net.offsets.items():
                if o < net.n_trainable and w[name][j].grad is not None:
                    gref[o:o + n] = w[name][j].grad.numpy().reshape(-1)
            l0, g0 = first_grad(up, tm, x, hr)
            up2, tm2 = build(rbn, lr, dtype, gw, vw)
            _, g1 = first_grad(up2, tm2, x, hr)
            up3, tm3 = build(rbn, lr, dtype, gw, vw)
            _, g2 = first_grad(up3, tm3, x[::-1].copy(), hr[::-1].copy())
            print("== %d residual blocks, %d patches %dx%d, storage %s, gradient scale %s" % (
                rbn, B, lr, lr, str(dtype).replace("torch.", ""), gs if gs is not None else 1))
            print("    first loss %.6f vs oracle %.6f (rel %.2e); non-finite gradient elements: %d" % (
                l0, ref_losses[0], abs(l0 - ref_losses[0]) / ref_losses[0], int(np.sum(~np.isfinite(g0)))))
            summarize("run-to-run (same data)", per_tensor(net, g1, g0))
            summarize("reversed batch (floor)", per_tensor(net, g2, g0))
            summarize("vs fp32 oracle", per_tensor(net, g0, gref))
            up4, tm4 = build(rbn, lr, dtype, gw, vw)
            losses = [tm4.train_on_batch(*data(B, lr, s)) for s in range(4)]
            print("    4 free-running steps: ours %s oracle %s  max rel diff %.4f" % (
                np.round(losses, 5), np.round(ref_losses, 5), float(np.max(np.abs(np.array(losses) - ref_losses) / np.array(ref_losses)))))
            sys.stdout.flush()


if __name__ == "__main__":
    main()
