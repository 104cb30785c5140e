"""-m gpu: BASELINE.json configs[0] on the reference's own frame, and the saturation guard of the fp16 inference storage.

example_movie_frames/example_frame_test.ipynb cells 2, 4, 9, 10: BtoA_00286.jpg -> RGB -> crop (0, 0, 480, 270) ->
np.array -> swapaxes(1, 2) (the notebook feeds WIDTH-first frames: (1, 480, 270, 3)) -> x / 127.5 - 1 ->
upscaler.predict.  The decoded crop is the committed fixture tests/golden/BtoA_00286_crop480x270.png
(tools/make_real_frame_fixture.py).  The trained model_upscaler_orig_vgg-mse_best.h5 is a missing blob of the reference
checkout, so the weights are the seeded synthetic set of SURVEY.md 8(d); what this adds over the U(-1,1) noise frames of
the other tests is natural-image statistics: flat regions, edges, PReLU dead zones, tanh tails.
Tolerance: north_star's max-abs <= 2e-2 on [-1,1]; PSNR >= 50 dB against the fp32 oracle.
"""
import os
import warnings

import numpy as np
import pytest
import torch

from oracle import srgan_oracle as O

pytestmark = pytest.mark.gpu
FIXTURE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "BtoA_00286_crop480x270.png")


def _frame(width_first):
    from PIL import Image
    from upscaler.data import convert_image_series_to_array
    a = np.array(Image.open(FIXTURE).convert("RGB"))
    assert a.shape == (270, 480, 3)
    arr = np.array([a])
    if width_first:
        arr = arr.swapaxes(1, 2)                                   # cell 2: img_array.swapaxes(1, 2)
    ref_arr = (arr.astype(np.float32) / 127.5) - 1                 # cell 2
    ours = convert_image_series_to_array([a.swapaxes(0, 1) if width_first else a])
    assert np.allclose(ours, ref_arr, atol=1e-6)
    return np.ascontiguousarray(ref_arr.astype(np.float32))


def _psnr(a, b):
    return 10 * np.log10(4.0 / max(float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)), 1e-20))


@pytest.mark.parametrize("ks", [3, 5])
@pytest.mark.parametrize("width_first", [True, False])
def test_config1_real_frame_matches_oracle(width_first, ks):
    """Model.predict on the notebook's frame, both orientations, k = 3 (shipped-weights architecture) and k = 5
    (model.py default): fp16 storage against the fp32 oracle."""
    from upscaler.model import make_upscaler_orig
    x = _frame(width_first)
    w = O.init_upscaler_orig_weights(seed=1234, kernel_size=ks)
    with torch.no_grad():
        ref = O.upscaler_orig_forward(O.to_torch(w), x, training=False).numpy()
    up = make_upscaler_orig((4 * x.shape[1], 4 * x.shape[2], 3), kernel_size=ks)
    up.net.set_weights(w)
    with warnings.catch_warnings():
        warnings.simplefilter("error")                             # a saturation fallback would be a failure here
        y = up.predict(x)
    assert y.shape == ref.shape == (1, 4 * x.shape[1], 4 * x.shape[2], 3)
    err, psnr = float(np.abs(y - ref).max()), _psnr(y, ref)
    u8_diff = int(np.abs(O.convert_array_to_u8(y).astype(np.int32) - O.convert_array_to_u8(ref).astype(np.int32)).max())
    print("real frame %s k=%d: max-abs %.5f, PSNR %.1f dB, max uint8 difference %d; saturation report %s" % (
        "480x270" if width_first else "270x480", ks, err, psnr, u8_diff, up.net.saturation_report))
    assert err <= 2e-2 and psnr >= 50.0
    assert up.net.saturation_report == {}
    assert u8_diff <= 3                                            # 2e-2 on [-1, 1] is 2.55 grey levels


def test_saturating_activations_are_reported_not_clipped():
    """The conv epilogues store fp16 with a saturating conversion.  With weights scaled so that the residual trunk leaves
    the fp16 range, the first forward pass must say so: 'raise' -> FloatingPointError naming the layer; 'fallback' ->
    RuntimeWarning, bf16 storage from then on, and a result that matches the oracle at bf16 accuracy instead of a
    silently clipped one."""
    from upscaler.networks import UpscalerOrigNet
    w = O.init_upscaler_orig_weights(seed=7, kernel_size=3, res_block_num=4)
    # head activations of std ~1e5 (about half of them beyond 65504) ride the BatchNorm-free skip stream into the first
    # Conv2DTranspose, whose weights are scaled back down so that the output stays inside tanh's active range
    w["initial/conv"][0] = w["initial/conv"][0] * 6.0e5
    w["upscaling/0/block/conv_transp"][0] = w["upscaling/0/block/conv_transp"][0] / 6.0e5
    x = np.random.default_rng(0).uniform(-1, 1, (1, 24, 32, 3)).astype(np.float32)
    with torch.no_grad():
        ref = O.upscaler_orig_forward(O.to_torch(w), x, training=False, res_block_num=4).numpy()
    xd = torch.from_numpy(x).cuda()
    net = UpscalerOrigNet(kernel_size=3, res_block_num=4, weights=w)
    net.on_saturation = "raise"
    with pytest.raises(FloatingPointError) as ei:
        net.forward_infer(xd)
    assert "initial/conv" in str(ei.value) and net.saturation_report.get("initial/conv", 0) > 0
    print("raise mode: %s" % ei.value)
    net2 = UpscalerOrigNet(kernel_size=3, res_block_num=4, weights=w)
    with pytest.warns(RuntimeWarning, match="bfloat16"):
        y = net2.forward_infer(xd)
    torch.cuda.synchronize()
    assert net2.infer_dtype == torch.bfloat16 and net2.saturation_report == {}
    y = y.cpu().numpy()
    clipped = UpscalerOrigNet(kernel_size=3, res_block_num=4, weights=w)
    yc = clipped.forward_infer(xd, check_saturation=False).cpu().numpy()      # what a silent clamp would have returned
    e_fb, e_clip = float(np.abs(y - ref).mean()), float(np.abs(yc - ref).mean())
    print("mean-abs error vs oracle: bf16 fallback %.4f, silently clipped fp16 %.4f" % (e_fb, e_clip))
    assert e_fb < 0.5 * e_clip
    # a second pass with unchanged weights does not pay for the check again, a weight change re-arms it
    assert net2._sat_pending is False
    net2.set_weights(w)
    assert net2._sat_pending is True


def test_count_saturated_kernel():
    from upscaler import ops
    for dt, big in ((torch.float16, 65504.0), (torch.bfloat16, float("inf"))):
        x = torch.randn(100003, device="cuda").to(dt)
        idx = torch.tensor([0, 7, 8, 4097, 100000, 100002], device="cuda")
        x[idx[:3]] = big
        x[idx[3:5]] = -big
        x[idx[5]] = float("nan")
        c = torch.zeros(1, dtype=torch.int32, device="cuda")
        ops.count_saturated(x, c)
        ops.count_saturated(x[:8], c)
        assert int(c.item()) == 6 + 2
