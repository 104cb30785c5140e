"""-m gpu parity of make_upscaler_incep_resnet (reference upscaling/upscaler/model.py:445-497 with inception_mini_resblock
:372-382, inception_resblock_3path :389-413, inception_resblock_2path :420-443) against the fp32 CPU
oracle's literal graph (BatchNorm -> PReLU -> Conv2D mini blocks, Concatenate, 1x1 conv, Add): the B200 path pads every path
to 64 channels, folds each BatchNorm + PReLU into the preceding conv's epilogue and replaces Concatenate + 1x1 conv by a chain
of per-path 1x1 convs through the residual operand, so this checks all of that at once."""
import numpy as np
import pytest
import torch

from oracle import srgan_oracle as O

pytestmark = pytest.mark.gpu


def _psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 10 * np.log10(4.0 / max(mse, 1e-20))


@pytest.mark.parametrize("shape", [(1, 24, 40), (2, 17, 23), (3, 9, 11)])
def test_incep_resnet_generator_inference_matches_oracle(shape):
    from upscaler.networks import UpscalerIncepResnetNet
    B, H, W = shape
    w = O.init_upscaler_incep_resnet_weights(seed=41)
    x = np.random.default_rng(13).uniform(-1, 1, (B, H, W, 3)).astype(np.float32)
    with torch.no_grad():
        ref = O.upscaler_incep_resnet_forward(O.to_torch(w), x).numpy()
    net = UpscalerIncepResnetNet(weights=w)
    y = net.forward_infer(torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    y = y.cpu().numpy()
    err, psnr = float(np.abs(y - ref).max()), _psnr(y, ref)
    print("incep-resnet %s: max-abs %.5f, PSNR %.1f dB (output |mean| %.3f)" % (shape, err, psnr, float(np.abs(ref).mean())))
    assert y.shape == ref.shape == (B, 4 * H, 4 * W, 3)
    assert err <= 2e-2 and psnr >= 50.0


def test_incep_resnet_block_variants_and_kernel_sizes():
    """Non-default groups: 2-path blocks with a 5-wide 1xk / kx1 pair first, 3-path blocks with k = 5, no B group."""
    from upscaler.networks import UpscalerIncepResnetNet
    cfg = dict(a=("2path", 2, 5), b=("3path", 2, 5), c=("3path", 0, 3))
    w = O.init_upscaler_incep_resnet_weights(seed=42, **cfg)
    x = np.random.default_rng(14).uniform(-1, 1, (1, 20, 28, 3)).astype(np.float32)
    with torch.no_grad():
        ref = O.upscaler_incep_resnet_forward(O.to_torch(w), x, **cfg).numpy()
    net = UpscalerIncepResnetNet(weights=w, **cfg)
    y = net.forward_infer(torch.from_numpy(x).cuda()).cpu().numpy()
    err, psnr = float(np.abs(y - ref).max()), _psnr(y, ref)
    print("incep-resnet variants: max-abs %.5f, PSNR %.1f dB" % (err, psnr))
    assert err <= 2e-2 and psnr >= 50.0


def test_incep_resnet_full_frame_270x480():
    from upscaler.model import make_upscaler_incep_resnet
    w = O.init_upscaler_incep_resnet_weights(seed=43)
    x = np.random.default_rng(15).uniform(-1, 1, (1, 270, 480, 3)).astype(np.float32)
    with torch.no_grad():
        ref = O.upscaler_incep_resnet_forward(O.to_torch(w), x).numpy()
    m = make_upscaler_incep_resnet((1080, 1920, 3))
    m.net.set_weights(w)
    y = m.predict(x)
    err, psnr = float(np.abs(y - ref).max()), _psnr(y, ref)
    print("incep-resnet full frame: max-abs %.5f, PSNR %.1f dB" % (err, psnr))
    assert y.shape == (1, 1080, 1920, 3) and err <= 2e-2 and psnr >= 50.0


def test_incep_resnet_model_surface_save_load_and_compile_guard(tmp_path):
    from upscaler.model import VGG_LOSS, compile_training_model, load_model, make_upscaler_incep_resnet
    kw = dict(a_block_num=2, b_block_num=1, c_block_num=1)
    m = make_upscaler_incep_resnet((64, 96, 3), **kw)
    w = O.init_upscaler_incep_resnet_weights(seed=44, a=("3path", 2, 3), b=("2path", 1, 7), c=("2path", 1, 3))
    m.net.set_weights(w)
    assert m.input_shape == (None, 16, 24, 3) and m.output_shape == (None, 64, 96, 3)
    assert m.count_params() == O.count_params(w)
    x = np.random.default_rng(16).uniform(-1, 1, (2, 16, 24, 3)).astype(np.float32)
    y = m.predict(x)
    path = str(tmp_path / "model_upscaler_incep.h5")
    m.save(path)
    m2 = load_model(path)
    assert type(m2.net).__name__ == "UpscalerIncepResnetNet" and m2.net.block_cfg == m.net.block_cfg
    assert np.array_equal(m2.predict(x), y)
    with pytest.raises(NotImplementedError):
        compile_training_model(m, VGG_LOSS((64, 96, 3)).loss)
