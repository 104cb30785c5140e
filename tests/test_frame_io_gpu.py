"""-m gpu: the frame I/O edges (data.py:253-270) on the device are bit-exact with the reference's NumPy expressions, and
Model.predict_uint8 equals convert_image_series_to_array -> predict -> convert_array_to_image's arithmetic."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_u8_to_f32_norm_is_bit_exact():
    from upscaler import ops
    rng = np.random.default_rng(0)
    for n in (256, 16 * 1000 + 7, 3 * 270 * 480):
        x = rng.integers(0, 256, n, dtype=np.uint8)
        x[:256 if n >= 256 else n] = np.arange(256, dtype=np.uint8)[:n]
        ref = ((x / 127.5) - 1).astype(np.float32)          # data.py:266-270, then Keras' cast to float32
        got = ops.u8_to_f32_norm(torch.from_numpy(x).cuda()).cpu().numpy()
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_f32_to_u8_denorm_is_bit_exact():
    from upscaler import ops
    rng = np.random.default_rng(1)
    a = rng.uniform(-1, 1, 1 << 20).astype(np.float32)
    k = np.arange(0, 255, dtype=np.float64) + 0.5            # inputs that land on or next to the rounding ties
    ties = (k / 127.5 - 1).astype(np.float32)
    a[:ties.size] = ties
    a[ties.size:ties.size + 4] = [-1.0, 1.0, 0.0, -0.0]
    a = np.concatenate([a, np.nextafter(ties, np.float32(2)), np.nextafter(ties, np.float32(-2)), a[:5]])
    ref = np.uint8(np.around((a + 1) * 127.5))               # data.py:253-256 on a float32 prediction
    got = ops.f32_to_u8_denorm(torch.from_numpy(a).cuda()).cpu().numpy()
    assert np.array_equal(got, ref)


def test_predict_uint8_equals_numpy_conversions_around_predict():
    from upscaler.data import convert_array_to_uint8, convert_image_series_to_array, upscale_images
    from upscaler.model import make_upscaler_orig
    rng = np.random.default_rng(2)
    frames = rng.integers(0, 256, (5, 24, 40, 3), dtype=np.uint8)
    up = make_upscaler_orig((96, 160, 3), kernel_size=3, res_block_num=2)
    x = convert_image_series_to_array(list(frames))
    ref = convert_array_to_uint8(up.predict(x, batch_size=4))
    got = up.predict_uint8(frames, batch_size=4)
    assert got.dtype == np.uint8 and got.shape == (5, 96, 160, 3)
    assert np.array_equal(got, ref)
    imgs = upscale_images([f for f in frames], up)
    assert len(imgs) == 5 and imgs[0].size == (160, 96) and np.array_equal(np.asarray(imgs[3]), ref[3])
    with pytest.raises(ValueError):
        up.predict_uint8(frames.astype(np.float32))
