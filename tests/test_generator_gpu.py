"""-m gpu parity: generator inference through the CUDA path (C ABI) vs the fp32 CPU oracle.

Tolerance (BASELINE.json north_star): max-abs error <= 2e-2 on [-1,1] outputs and PSNR delta <= 0.05 dB.
With synthetic data there is no ground-truth image, so the PSNR bar is applied as PSNR(ours, oracle) >= 50 dB: noise
at 50 dB added to a ~30 dB reconstruction moves its PSNR by < 0.05 dB.

The inference path stores activations / weights as fp16 (same tcgen05 kind::f16 rate as bf16).  bf16 storage is kept
as an option and checked against the rounding floor measured by CPU emulation of bf16 storage: with the
SURVEY 8(d) weight distributions the trunk reaches std ~16 and bf16 rounding alone gives max-abs ~0.09 / 44 dB.
"""
import numpy as np
import pytest
import torch

from oracle import srgan_oracle as O

pytestmark = pytest.mark.gpu


def _psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 10 * np.log10(4.0 / max(mse, 1e-20))


def _run(ks, shape, dtype, seed=1234):
    from upscaler.networks import UpscalerOrigNet
    B, H, W = shape
    w = O.init_upscaler_orig_weights(seed=seed, kernel_size=ks)
    rng = np.random.default_rng(0)
    x = rng.uniform(-1, 1, (B, H, W, 3)).astype(np.float32)
    with torch.no_grad():
        ref = O.upscaler_orig_forward(O.to_torch(w), x, training=False).numpy()
    net = UpscalerOrigNet(kernel_size=ks, weights=w, infer_dtype=dtype)
    y = net.forward_infer(torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    y = y.cpu().numpy()
    assert y.shape == (B, 4 * H, 4 * W, 3)
    assert np.isfinite(y).all()
    return float(np.abs(y - ref).max()), _psnr(y, ref)


@pytest.mark.parametrize("ks,shape", [(3, (1, 24, 40)), (5, (2, 17, 23)), (3, (1, 40, 24)), (3, (3, 9, 11))])
def test_generator_inference_matches_oracle(ks, shape):
    err, psnr = _run(ks, shape, torch.float16)
    print("fp16 ks=%d %s max_abs=%.5f psnr=%.2f dB" % (ks, shape, err, psnr))
    assert err <= 2e-2          # north_star: max-abs <= 2e-2 on [-1,1]
    assert psnr >= 50.0         # => PSNR delta <= 0.05 dB on a ~30 dB reconstruction


@pytest.mark.parametrize("ks,shape", [(3, (1, 24, 40)), (5, (1, 17, 23))])
def test_generator_inference_bf16_storage_at_rounding_floor(ks, shape):
    err, psnr = _run(ks, shape, torch.bfloat16)
    print("bf16 ks=%d %s max_abs=%.5f psnr=%.2f dB" % (ks, shape, err, psnr))
    assert err <= 0.15 and psnr >= 42.0  # bf16 storage floor for these weights (CPU emulation: 0.0875 / 44.3 dB)


@pytest.mark.parametrize("ks,shape", [(3, (1, 24, 40)), (5, (2, 17, 23)), (3, (1, 40, 33))])
def test_skip_con_generator_inference_matches_oracle(ks, shape):
    """make_upscaler_skip_con (model.py:332-363): 224-filter ConvT blocks padded to 256, bilinear-resized input written
    into padding channels, permuted final-conv weights -- against the oracle's literal Concatenate graph."""
    from upscaler.networks import UpscalerSkipConNet
    B, H, W = shape
    w = O.init_upscaler_skip_con_weights(seed=21, kernel_size=ks)
    x = np.random.default_rng(5).uniform(-1, 1, (B, H, W, 3)).astype(np.float32)
    with torch.no_grad():
        ref = O.upscaler_skip_con_forward(O.to_torch(w), x).numpy()
    net = UpscalerSkipConNet(kernel_size=ks, weights=w)
    y = net.forward_infer(torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    y = y.cpu().numpy()
    err, psnr = float(np.abs(y - ref).max()), _psnr(y, ref)
    print("skip-con k=%d %s: max-abs %.5f, PSNR %.1f dB" % (ks, shape, err, psnr))
    assert y.shape == ref.shape and err <= 2e-2 and psnr >= 50.0


def test_resize_bilinear_into_matches_oracle():
    from upscaler import ops
    x = np.random.default_rng(6).uniform(-1, 1, (2, 9, 13, 3)).astype(np.float32)
    out = torch.zeros((2, 36, 52, 64), dtype=torch.float16, device="cuda")
    ops.resize_bilinear_into(torch.from_numpy(x).cuda(), out, 24, 4)
    torch.cuda.synchronize()
    ref = O.resize_bilinear_tf1(torch.from_numpy(x).permute(0, 3, 1, 2), 4).permute(0, 2, 3, 1).numpy()
    got = out.float().cpu().numpy()
    assert np.abs(got[..., 24:27] - ref).max() < 1e-3          # fp16 storage of values in [-1, 1]
    assert np.all(got[..., :24] == 0) and np.all(got[..., 27:] == 0)
