"""-m gpu parity of make_upscaler_attention (reference upscaling/upscaler/model.py:296-328 with residual_block_attention
:30-48 and upsampling_block_attention :78-98) against the fp32 CPU oracle: the three memory-bound
kernels it adds, the whole generator (north_star bar: max-abs <= 2e-2, PSNR >= 50 dB), the Keras-HDF5 round trip and the
builder surface."""
import numpy as np
import pytest
import torch

from oracle import srgan_oracle as O

pytestmark = pytest.mark.gpu


def _psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 10 * np.log10(4.0 / max(mse, 1e-20))


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_gate_mul_kernel(dtype):
    from upscaler import ops
    rng = np.random.default_rng(1)
    att = torch.from_numpy(rng.standard_normal((2, 7, 9, 256)).astype(np.float32) * 3).to(dtype).cuda()
    m = torch.from_numpy(rng.standard_normal((2, 7, 9, 64)).astype(np.float32) * 10).to(dtype).cuda()
    out = ops.gate_mul(att, m, att_c0=128)
    torch.cuda.synchronize()
    want = (torch.sigmoid(att[..., 128:192].float()) * m.float()).cpu().numpy()
    got = out.float().cpu().numpy()
    assert np.abs(got - want).max() <= (2e-3 if dtype == torch.float16 else 1.6e-2) * np.abs(want).max()
    with pytest.raises(Exception):
        ops.gate_mul(att, m, att_c0=200)            # slice [200, 264) leaves the 256 attention channels


@pytest.mark.parametrize("factor", [1, 2, 3])
def test_resize_nearest_bilinear_pair_kernel(factor):
    from upscaler import ops
    x = np.random.default_rng(2).uniform(-1, 1, (2, 5, 7, 3)).astype(np.float32)
    got = ops.resize_nearest_bilinear_pair(torch.from_numpy(x).cuda(), factor).cpu().numpy()
    xt = torch.from_numpy(x).permute(0, 3, 1, 2)
    near = O.resize_nearest_tf1(xt, factor).permute(0, 2, 3, 1).numpy()
    bil = O.resize_bilinear_tf1(xt, factor).permute(0, 2, 3, 1).numpy()
    assert got.shape == (2, 5 * factor, 7 * factor, 6)
    assert np.array_equal(got[..., :3], near)                  # a copy: bit-exact
    assert np.abs(got[..., 3:] - bil).max() < 1e-6


@pytest.mark.parametrize("stride", [2, 4])
def test_atanh_conv_transpose_kernel(stride):
    """Conv2DTranspose(128, stride + 1, strides=stride, 'same')(atanh(0.99999 x)) including saturated pixels x = +-1."""
    from upscaler import ops
    rng = np.random.default_rng(3)
    x = rng.uniform(-1, 1, (2, 6, 5, 3)).astype(np.float32)
    x[0, 0, 0] = 1.0
    x[1, 5, 4] = -1.0
    k = stride + 1
    w = (rng.standard_normal((k, k, 128, 3)) * 0.2).astype(np.float32)
    b = rng.standard_normal(128).astype(np.float32)
    got = ops.atanh_conv2d_transpose_small(torch.from_numpy(x).cuda(), torch.from_numpy(w).cuda(), torch.from_numpy(b).cuda(),
                                           stride, dtype=torch.float16).float().cpu().numpy()
    xt = torch.atanh(0.99999 * torch.from_numpy(x).permute(0, 3, 1, 2))
    want = O.conv2d_transpose_same(xt, torch.from_numpy(w), torch.from_numpy(b), stride).permute(0, 2, 3, 1).numpy()
    assert got.shape == want.shape == (2, 6 * stride, 5 * stride, 128)
    assert np.abs(got - want).max() <= 2e-3 * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("ks,shape", [(3, (1, 24, 40)), (5, (2, 17, 23)), (3, (1, 40, 33)), (3, (3, 9, 11))])
def test_attention_generator_inference_matches_oracle(ks, shape):
    from upscaler.networks import UpscalerAttentionNet
    B, H, W = shape
    w = O.init_upscaler_attention_weights(seed=31, kernel_size=ks)
    x = np.random.default_rng(7).uniform(-1, 1, (B, H, W, 3)).astype(np.float32)
    x[0, 0, 0] = (1.0, -1.0, 1.0)                # saturated pixels: the atanh(0.99999 x) branch at its steepest
    with torch.no_grad():
        ref = O.upscaler_attention_forward(O.to_torch(w), x).numpy()
    net = UpscalerAttentionNet(kernel_size=ks, weights=w)
    y = net.forward_infer(torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    y = y.cpu().numpy()
    err, psnr = float(np.abs(y - ref).max()), _psnr(y, ref)
    print("attention k=%d %s: max-abs %.5f, PSNR %.1f dB" % (ks, shape, err, psnr))
    assert y.shape == ref.shape == (B, 4 * H, 4 * W, 3)
    assert err <= 2e-2 and psnr >= 50.0


def test_attention_generator_full_frame_270x480():
    """configs[1]'s frame through make_upscaler_attention (k = 3): production tilings of every kernel it uses."""
    from upscaler.model import make_upscaler_attention
    w = O.init_upscaler_attention_weights(seed=32, kernel_size=3)
    x = np.random.default_rng(8).uniform(-1, 1, (1, 270, 480, 3)).astype(np.float32)
    with torch.no_grad():
        ref = O.upscaler_attention_forward(O.to_torch(w), x).numpy()
    m = make_upscaler_attention((1080, 1920, 3), kernel_size=3)
    m.net.set_weights(w)
    y = m.predict(x)
    err, psnr = float(np.abs(y - ref).max()), _psnr(y, ref)
    print("attention full frame: max-abs %.5f, PSNR %.1f dB" % (err, psnr))
    assert y.shape == (1, 1080, 1920, 3) and err <= 2e-2 and psnr >= 50.0


def test_attention_model_surface_save_load_and_compile_guard(tmp_path):
    from upscaler.model import VGG_LOSS, compile_training_model, load_model, make_upscaler_attention
    m = make_upscaler_attention((64, 96, 3), kernel_size=3, res_block_num=2)
    assert m.input_shape == (None, 16, 24, 3) and m.output_shape == (None, 64, 96, 3)
    w = O.init_upscaler_attention_weights(seed=33, kernel_size=3, res_block_num=2)
    m.net.set_weights(w)
    assert m.count_params() == O.count_params(w)
    x = np.random.default_rng(9).uniform(-1, 1, (2, 16, 24, 3)).astype(np.float32)
    y = m.predict(x)
    path = str(tmp_path / "model_upscaler_att.h5")
    m.save(path)
    m2 = load_model(path)
    assert type(m2.net).__name__ == "UpscalerAttentionNet" and m2.net.res_block_num == 2
    assert np.array_equal(m2.predict(x), y)
    with pytest.raises(NotImplementedError):
        compile_training_model(m, VGG_LOSS((64, 96, 3)).loss)
