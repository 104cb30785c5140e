"""-m gpu edge cases of the public API (Model.predict / predict_uint8 / train_on_batch): empty and ragged batches, frames
smaller than any kernel tile, odd sizes, shape errors (Keras raises ValueError), and a frame larger than the headline size."""
import numpy as np
import pytest
import torch

from oracle import srgan_oracle as O

pytestmark = pytest.mark.gpu


def _model(ks=3, hw=(8, 12)):
    from upscaler.model import make_upscaler_orig
    m = make_upscaler_orig((4 * hw[0], 4 * hw[1], 3), kernel_size=ks)
    m.net.set_weights(O.init_upscaler_orig_weights(seed=2, kernel_size=ks))
    return m


def test_predict_empty_and_ragged_batches():
    m = _model()
    x = np.random.default_rng(0).uniform(-1, 1, (5, 8, 12, 3)).astype(np.float32)
    y0 = m.predict(x[:0])
    assert y0.shape == (0, 32, 48, 3) and y0.dtype == np.float32
    y_all = m.predict(x, batch_size=8)                  # batch_size larger than the batch
    y_rag = m.predict(x, batch_size=2)                  # 2 + 2 + 1
    y_one = np.concatenate([m.predict(x[i:i + 1]) for i in range(5)])
    assert y_all.shape == (5, 32, 48, 3)
    assert np.array_equal(y_all, y_rag)
    # one frame per call: for frames this small the tiling (pixels of several frames share an MMA tile) and with it the
    # fp32 summation order depend on the batch, so 16-bit roundings can differ -- equal to rounding noise, not bit for bit
    # (at 270x480 frames are bit-identical whatever the batch: tests/test_fullsize_gpu.py)
    assert np.abs(y_all - y_one).max() < 1.5e-2        # seen: 6e-3; each side is within 2e-2 of the oracle
    u = np.random.default_rng(1).integers(0, 256, (3, 8, 12, 3), dtype=np.uint8)
    assert m.predict_uint8(u[:0]).shape == (0, 32, 48, 3)
    assert np.array_equal(m.predict_uint8(u, batch_size=2), m.predict_uint8(u, batch_size=8))


@pytest.mark.parametrize("hw", [(1, 1), (1, 7), (3, 2), (2, 33)])
def test_tiny_and_thin_frames_match_oracle(hw):
    from upscaler.networks import UpscalerOrigNet
    w = O.init_upscaler_orig_weights(seed=2, kernel_size=3)
    x = np.random.default_rng(hw[0] * 100 + hw[1]).uniform(-1, 1, (2, hw[0], hw[1], 3)).astype(np.float32)
    with torch.no_grad():
        ref = O.upscaler_orig_forward(O.to_torch(w), x).numpy()
    y = UpscalerOrigNet(kernel_size=3, weights=w).forward_infer(torch.from_numpy(x).cuda()).cpu().numpy()
    assert y.shape == ref.shape and np.abs(y - ref).max() <= 2e-2


def test_shape_errors_are_value_errors():
    m = _model()
    with pytest.raises(ValueError):
        m.predict(np.zeros((1, 9, 12, 3), np.float32))        # wrong height for a model bound to 8 x 12
    with pytest.raises(ValueError):
        m.predict(np.zeros((1, 8, 12, 4), np.float32))        # wrong channel count
    with pytest.raises((ValueError, TypeError)):
        m.net.forward_infer(torch.zeros((1, 8, 12, 3), dtype=torch.float16, device="cuda"))


def test_frame_larger_than_the_headline_size():
    """540 x 960 -> 2160 x 3840 (a 4K frame: 4x the pixels of configs[1]); properties only -- finite, in tanh range, and
    the top-left quadrant equals the same quadrant pushed through alone away from the cut."""
    from upscaler.networks import UpscalerOrigNet
    net = UpscalerOrigNet(kernel_size=3, weights=O.init_upscaler_orig_weights(seed=1234, kernel_size=3))
    x = torch.from_numpy(np.random.default_rng(9).uniform(-1, 1, (1, 540, 960, 3)).astype(np.float32)).cuda()
    y = net.forward_infer(x).clone()
    assert tuple(y.shape) == (1, 2160, 3840, 3) and bool(torch.isfinite(y).all()) and float(y.abs().max()) <= 1.0
    q = net.forward_infer(x[:, :270, :480].contiguous())
    torch.cuda.synchronize()
    m = 48
    assert torch.equal(q[0, :4 * (270 - m), :4 * (480 - m)], y[0, :4 * (270 - m), :4 * (480 - m)])


def test_train_on_batch_with_one_sample_and_odd_patch():
    """A batch of one (BatchNorm statistics over the pixels of a single sample) and a patch size that is not a multiple
    of any tile (LR 20 x 28)."""
    from upscaler.keras_compat import Adam
    from upscaler.model import VGG_MSE_LOSS, compile_training_model, make_upscaler_orig
    gw, vw = O.init_upscaler_orig_weights(seed=11, kernel_size=3, res_block_num=2), O.init_vgg19_weights()
    rng = np.random.default_rng(3)
    hr = rng.uniform(-1, 1, (1, 80, 112, 3)).astype(np.float32)
    x = hr.reshape(1, 20, 4, 28, 4, 3).mean(axis=(2, 4)).astype(np.float32)
    up = make_upscaler_orig((80, 112, 3), kernel_size=3, res_block_num=2)
    up.net.set_weights(gw)
    tm = compile_training_model(up, VGG_MSE_LOSS((80, 112, 3), 1e-3, vgg19=vw).loss, optimizer=Adam())
    got = tm.train_on_batch(x, hr)
    want = O.GeneratorTrainer(gw, vw, "vgg-mse", 1e-3, res_block_num=2).train_on_batch(x, hr)
    assert abs(got - want) <= 1e-2 * want


def test_train_on_batch_takes_float64_and_strided_batches():
    """The reference's loaders hand train_on_batch float64 arrays (data.py:266-270: array / 127.5 - 1) and slices of larger
    arrays; the staging copy casts them to float32 on its way into pinned memory.  Same weights: same loss as the float32,
    contiguous batch (large enough for the multi-threaded staging path)."""
    from upscaler.keras_compat import Adam
    from upscaler.model import VGG_MSE_LOSS, compile_training_model, make_upscaler_orig
    gw, vw = O.init_upscaler_orig_weights(seed=4, kernel_size=3, res_block_num=1), O.init_vgg19_weights()
    rng = np.random.default_rng(12)
    hr = rng.uniform(-1, 1, (4, 192, 192, 3)).astype(np.float32)
    x = hr.reshape(4, 48, 4, 48, 4, 3).mean(axis=(2, 4)).astype(np.float32)
    losses = []
    for variant in range(3):
        up = make_upscaler_orig((192, 192, 3), kernel_size=3, res_block_num=1)
        up.net.set_weights(gw)
        tm = compile_training_model(up, VGG_MSE_LOSS((192, 192, 3), 1e-3, vgg19=vw).loss, optimizer=Adam())
        if variant == 0:
            xa, ya = x, hr
        elif variant == 1:
            xa, ya = x.astype(np.float64), hr.astype(np.float64)
        else:                                                        # every other sample of a twice larger array
            xa, ya = np.repeat(x, 2, axis=0)[::2], np.repeat(hr, 2, axis=0)[::2]
            assert not ya.flags.c_contiguous
        losses.append(tm.train_on_batch(xa, ya))
    # equal up to the summation order of the loss reduction's atomics (run-to-run: ~5e-7 relative)
    assert max(losses) - min(losses) <= 1e-5 * abs(losses[0]), losses
