"""CPU tests pinning the oracle: shape / parameter-count known answers recorded in the reference's notebooks
(upscaling/cnn_test.ipynb cells 12 and 18), TensorFlow padding rules against a pure-Python loop restatement, Keras
BatchNorm / Adam / loss arithmetic by hand-computed values."""
import math

import numpy as np
import torch

from oracle import srgan_oracle as O
from oracle import tf_loops as L


def test_same_padding_shapes_from_cnn_test_notebook():
    # cnn_test.ipynb cell 12: Conv2D k3 s2 on (135,240) -> (68,120); four Conv2DTranspose s2 -> 136x240 ... 1088x1920
    assert O.tf_same_pad(135, 3, 2)[2] == 68 and O.tf_same_pad(240, 3, 2)[2] == 120
    x = torch.zeros(1, 1, 68, 120)
    k = torch.zeros(3, 3, 1, 1)
    shapes = []
    for _ in range(4):
        x = O.conv2d_transpose_same_s2(x, k)
        shapes.append(tuple(x.shape[2:]))
    assert shapes == [(136, 240), (272, 480), (544, 960), (1088, 1920)]
    # even input, stride 2, k3: pad (0,1); odd input: (1,1)  
    assert O.tf_same_pad(32, 3, 2)[:2] == (0, 1) and O.tf_same_pad(27, 3, 2)[:2] == (1, 1)
    assert O.tf_same_pad(30, 5, 1)[:2] == (2, 2) and O.tf_same_pad(30, 9, 1)[:2] == (4, 4)


def test_param_counts_from_notebooks_and_survey():
    # cnn_test.ipynb cell 12: Conv2D(1, 3) on 3 channels = 28 params; Conv2DTranspose(1, 3) on 1 channel = 10
    assert 3 * 3 * 3 * 1 + 1 == 28 and 3 * 3 * 1 * 1 + 1 == 10
    # cell 18: 9x9x3->128 conv = 31232, PReLU(shared_axes=[1,2]) = 128, 3x3 128->128 = 147584
    assert 9 * 9 * 3 * 128 + 128 == 31232 and 3 * 3 * 128 * 128 + 128 == 147584
    g3 = O.init_upscaler_orig_weights(kernel_size=3)
    g5 = O.init_upscaler_orig_weights(kernel_size=5)
    assert g3["initial/prelu"][0].shape == (1, 1, 64)
    assert O.count_params(g3) == 2043779 and O.count_params(g5) == 5517187
    d = O.init_discriminator_simple_512_weights(input_shape=(512, 512, 3))
    assert O.count_params(d) == 15499329
    v = O.init_vgg19_weights()
    assert len(v) == 16 and v["block5_conv4"][0].shape == (3, 3, 512, 512)


def test_conv_same_matches_loop_restatement():
    rng = np.random.default_rng(0)
    for (H, W, k, s) in [(6, 7, 3, 1), (8, 8, 3, 2), (7, 5, 3, 2), (6, 6, 5, 1), (5, 9, 9, 1)]:
        x = rng.standard_normal((H, W, 2)).astype(np.float32)
        kern = rng.standard_normal((k, k, 2, 3)).astype(np.float32)
        ref = L.conv2d_same_loops(x, kern, s)
        got = O.conv2d_same(torch.from_numpy(x).permute(2, 0, 1)[None], torch.from_numpy(kern), None, s)[0]
        np.testing.assert_allclose(got.permute(1, 2, 0).numpy(), ref, atol=1e-4)


def test_conv_transpose_matches_loop_restatement():
    rng = np.random.default_rng(1)
    for (H, W, k) in [(4, 5, 3), (3, 3, 5), (5, 4, 4), (2, 6, 2)]:
        x = rng.standard_normal((H, W, 2)).astype(np.float32)
        kern = rng.standard_normal((k, k, 3, 2)).astype(np.float32)
        ref = L.conv2d_transpose_same_s2_loops(x, kern)
        got = O.conv2d_transpose_same_s2(torch.from_numpy(x).permute(2, 0, 1)[None], torch.from_numpy(kern))[0]
        assert tuple(got.shape[1:]) == (2 * H, 2 * W)
        np.testing.assert_allclose(got.permute(1, 2, 0).numpy(), ref, atol=1e-4)
    # SURVEY.md 5.8 tap alignment, k=3: even y=2j -> x[j]W0 + x[j-1]W2, odd -> x[j]W1
    x = np.zeros((3, 1, 1), np.float32); x[1] = 1
    kern = np.array([10., 20., 30.], np.float32).reshape(3, 1, 1, 1) * np.ones((3, 1, 1, 1), np.float32)
    y = L.conv2d_transpose_same_s2_loops(x, np.broadcast_to(kern, (3, 1, 1, 1)).copy())[:, 0, 0]
    assert y.tolist() == [0, 0, 10, 20, 30, 0]


def test_generator_graph_shapes_and_modes():
    w = O.to_torch(O.init_upscaler_orig_weights(kernel_size=3))
    x = np.random.default_rng(0).uniform(-1, 1, (2, 8, 12, 3)).astype(np.float32)
    with torch.no_grad():
        y = O.upscaler_orig_forward(w, x)
        stats = []
        yt = O.upscaler_orig_forward(w, x, training=True, stats_out=stats)
    assert tuple(y.shape) == (2, 32, 48, 3) and float(y.abs().max()) <= 1.0
    assert len(stats) == 33 and len(O.upscaler_bn_names()) == 33
    assert not torch.allclose(y, yt)  # batch statistics differ from the (randomised) moving statistics


def test_vgg_and_discriminator_shapes():
    v = O.to_torch(O.init_vgg19_weights())
    with torch.no_grad():
        f = O.vgg19_block5_conv4(v, np.zeros((1, 48, 64, 3), np.float32))
    assert tuple(f.shape) == (1, 3, 4, 512) and float(f.min()) >= 0  # 4 max-pools, post-ReLU
    d = O.to_torch(O.init_discriminator_simple_512_weights(input_shape=(64, 64, 3)))
    assert d["discriminator/final/Dense_1"][0].shape == (512, 1024)  # 64 -> 1x1x512 after 8 stride-2 convs
    with torch.no_grad():
        out = O.discriminator_simple_512_forward(d, np.random.default_rng(0).standard_normal((3, 64, 64, 3)).astype(np.float32))
    assert tuple(out.shape) == (3, 1)


def test_batchnorm_training_and_moving_update():
    x = torch.tensor([[1.0, 2.0], [3.0, 6.0], [5.0, 10.0]])  # (N=3, C=2)
    stats = []
    y = O.batch_norm(x, torch.ones(2), torch.zeros(2), None, None, True, stats)
    m, v, n = stats[0]
    assert torch.allclose(m, torch.tensor([3.0, 6.0])) and torch.allclose(v, torch.tensor([8 / 3, 32 / 3])) and n == 3
    assert torch.allclose(y[:, 0], (x[:, 0] - 3) / math.sqrt(8 / 3 + 1e-3), atol=1e-6)
    mm, mv = O.update_moving(torch.zeros(2), torch.ones(2), m, v, n)
    assert torch.allclose(mm, 0.01 * m)
    assert torch.allclose(mv, 0.99 + 0.01 * v * (3 / (3 - 1.001)))


def test_keras_adam_first_steps():
    p = torch.tensor([1.0, -2.0])
    opt = O.KerasAdam()
    p.grad = torch.tensor([0.5, -0.25])
    opt.apply([("p", p)])
    # t=1: m = 0.1 g, v = 0.001 g^2, lr_t = lr*sqrt(0.001)/0.1  =>  step = lr * g/|g| (up to eps)
    assert torch.allclose(p, torch.tensor([1.0 - 1e-3, -2.0 + 1e-3]), atol=1e-6)
    assert opt.iterations == 1
    p.grad = torch.tensor([0.5, -0.25])
    opt.apply([("p", p)])
    assert torch.allclose(p, torch.tensor([1.0 - 2e-3, -2.0 + 2e-3]), atol=2e-6)


def test_losses():
    a, b = torch.tensor([[1.0], [3.0]]), torch.tensor([[2.0], [-1.0]])
    assert float(O.wasserstein_loss(a, b)) == (2.0 - 3.0) / 2
    d, g = O.gan_losses("wasserstein", "none", a, b)
    assert float(d) == 2.0 - 0.5 and float(g) == 0.5
    d, g = O.gan_losses("relativistic", "bi-log", a, b)
    x = 1.5
    assert abs(float(d) - (x / (1 + x)) * math.log(x + 2)) < 1e-6 and abs(float(g) + float(d)) < 1e-6
    assert abs(float(O.loss_activation("log-sigm", torch.tensor(0.0))) - math.log(0.5)) < 1e-6
    t, p = torch.zeros(1, 4, 4, 3), torch.full((1, 4, 4, 3), 0.5)
    assert abs(float(O.content_loss("mse", None, t, p, 0)) - 0.25) < 1e-7
    assert abs(float(O.content_loss("mae", None, t, p, 0)) - 0.5) < 1e-7


def test_image_io_contract():
    # data.py:253-270: x/127.5-1 and uint8(around((a+1)*127.5)) round-trip every byte value
    u = np.arange(256, dtype=np.uint8).reshape(16, 16, 1).repeat(3, 2)
    a = O.convert_image_to_array(u)
    assert a.min() == -1.0 and a.max() == 1.0
    assert (O.convert_array_to_u8(a) == u).all()
    assert O.convert_array_to_u8(np.array([0.0])).tolist() == [128]  # np.around: 127.5 -> 128 (round half to even)


def test_trainers_run_and_decrease_loss():
    torch.manual_seed(0)
    g = O.init_upscaler_orig_weights(seed=3, kernel_size=3, res_block_num=1, randomize=False)
    tr = O.GeneratorTrainer(g, None, loss_kind="mse", res_block_num=1)
    rng = np.random.default_rng(0)
    y = rng.uniform(-1, 1, (2, 16, 16, 3)).astype(np.float32)
    x = y.reshape(2, 4, 4, 4, 4, 3).mean(axis=(2, 4))
    losses = [tr.train_on_batch(x, y) for _ in range(6)]
    assert losses[-1] < losses[0]
    assert tr.opt.iterations == 6


def test_resize_bilinear_tf1_known_answers_and_loops():
    """K.resize_images(..., 'bilinear') of make_upscaler_skip_con (model.py:357) = TF 1.x resize_bilinear without
    half-pixel centres: [0, 1] enlarged x2 is [0, 0.5, 1, 1] (the last sample clamps), not torch's [0, 0.25, 0.75, 1]."""
    import torch
    from oracle import tf_loops
    x = torch.tensor([[[[0.0, 1.0]]]])
    assert O.resize_bilinear_tf1(x, 2)[0, 0].tolist() == [[0.0, 0.5, 1.0, 1.0], [0.0, 0.5, 1.0, 1.0]]
    rng = np.random.default_rng(0)
    img = rng.uniform(-1, 1, (5, 7)).astype(np.float32)
    got = O.resize_bilinear_tf1(torch.from_numpy(img)[None, None], 4)[0, 0].numpy()
    want = np.array(tf_loops.resize_bilinear_tf1_loops(img.tolist(), 4), np.float32)
    assert got.shape == (20, 28) and np.abs(got - want).max() < 1e-6


def test_skip_con_oracle_shapes_and_param_count():
    w = O.init_upscaler_skip_con_weights(seed=3, kernel_size=3)
    # orig trunk (k=3) up to the after-res block, then 3x3x64x224 + 224, 3x3x224x224 + 224, 9x9x227x3 + 3
    trunk = O.count_params({k: v for k, v in O.init_upscaler_orig_weights(seed=3, kernel_size=3).items()
                            if not k.startswith("upscaling") and k != "final/conv"})
    assert O.count_params(w) == trunk + (9 * 64 * 224 + 224) + (9 * 224 * 224 + 224) + (81 * 227 * 3 + 3)
    x = np.random.default_rng(1).uniform(-1, 1, (1, 6, 5, 3)).astype(np.float32)
    import torch
    with torch.no_grad():
        y = O.upscaler_skip_con_forward(O.to_torch(w), x)
    assert tuple(y.shape) == (1, 24, 20, 3) and float(y.abs().max()) <= 1.0


def test_attention_generator_oracle_pieces_and_shapes():
    """make_upscaler_attention (model.py:296-328): the general-stride Conv2DTranspose and the nearest resize against the loop
    restatements of TensorFlow's rules; output shape; parameter count against the layer list of model.py:30-48, 78-98."""
    from oracle import tf_loops
    rng = np.random.default_rng(3)
    for k, s in ((3, 2), (5, 4), (5, 2), (4, 4)):
        x = rng.standard_normal((4, 5, 3)).astype(np.float32)
        ker = rng.standard_normal((k, k, 6, 3)).astype(np.float32)
        got = O.conv2d_transpose_same(torch.from_numpy(x).permute(2, 0, 1)[None], torch.from_numpy(ker), None, s)[0].permute(1, 2, 0).numpy()
        want = tf_loops.conv2d_transpose_same_loops(x, ker, s)
        assert got.shape == (4 * s, 5 * s, 6)
        assert np.abs(got - want).max() < 1e-4
    img = rng.standard_normal((3, 4)).astype(np.float32)
    near = O.resize_nearest_tf1(torch.from_numpy(img)[None, None], 2)[0, 0].numpy()
    assert np.array_equal(near, np.array(tf_loops.resize_nearest_tf1_loops(img.tolist(), 2), np.float32))
    for k in (3, 5):
        w = O.init_upscaler_attention_weights(seed=2, kernel_size=k)
        f, up = 64, 128
        want = (81 * 3 * f + f) + f                                          # initial conv + PReLU
        want += 16 * ((k * k * 3 * f + f) + 2 * (k * k * f * f + f) + 2 * 4 * f + f)      # attention conv, 2 convs, 2 BN, PReLU
        want += k * k * f * f + f + 4 * f                                     # after_res conv + BN
        for cin, scale in ((f, 2), (up, 4)):
            want += (k * k * 6 * cin + cin) + (k * k * cin * up + up) + ((scale + 1) ** 2 * 3 * up + up)
        want += 81 * up * 3 + 3
        assert O.count_params(w) == want
        with torch.no_grad():
            y = O.upscaler_attention_forward(O.to_torch(w), rng.uniform(-1, 1, (2, 6, 7, 3)).astype(np.float32))
        assert tuple(y.shape) == (2, 24, 28, 3) and float(y.abs().max()) <= 1.0


def test_incep_resnet_oracle_structure_matches_the_package_and_the_reference_layer_list():
    """make_upscaler_incep_resnet (model.py:445-497): filter counts int(64 * .3 / .4 / .5 / .75) = 19 / 25 / 32 / 48, layer names
    <block>/<path>/<i>/{batch_norm, prelu, KHxKW}; the oracle's weight keys equal the package's layer list (two independent
    restatements of the same graph), parameter counts agree, shapes propagate."""
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "video-cycle_gan-upscaling_b200"))
    from upscaler import networks
    w = O.init_upscaler_incep_resnet_weights(seed=5)
    specs = networks.upscaler_incep_resnet_layer_specs()
    assert [n for n, _, _ in specs] == list(w.keys())
    for name, kind, cfg in specs:
        for a, shp in zip(w[name], networks._shapes(kind, cfg)):
            assert tuple(a.shape) == tuple(shp), name
    assert tuple(w["inc_res_block/B/2p/3/b/2/1x7"][0].shape) == (1, 7, 19, 25)
    assert tuple(w["inc_res_block/B/2p/3/b/3/7x1"][0].shape) == (7, 1, 25, 32)
    assert tuple(w["inc_res_block/A/3p/0/c/2/3x3"][0].shape) == (3, 3, 32, 48)
    assert tuple(w["inc_res_block/A/3p/0/final/1x1"][0].shape) == (1, 1, 128, 64)
    assert tuple(w["inc_res_block/c/2p/4/final/1x1"][0].shape) == (1, 1, 64, 64)
    with torch.no_grad():
        y = O.upscaler_incep_resnet_forward(O.to_torch(w), np.random.default_rng(1).uniform(-1, 1, (2, 6, 7, 3)).astype(np.float32))
    assert tuple(y.shape) == (2, 24, 28, 3)
