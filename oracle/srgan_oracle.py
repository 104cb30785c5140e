"""CPU oracle for the SRGAN upscaler hot path — TEST INFRASTRUCTURE ONLY.

This is a PyTorch-CPU fp32 restatement of what the reference's Keras/TensorFlow graphs compute for the path named in
BASELINE.json (generator forward, generator/discriminator training step, VGG19 perceptual loss).  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / --impl reference legs may import it; the product package
never does.

PARITY UNPINNED: the reference cannot be executed in this environment (tensorflow / keras / h5py are absent, the
trained .h5 and the ImageNet VGG19 weights are missing blobs / downloads) and it ships no tests or value fixtures.
What IS pinned against the reference: layer shapes and parameter counts printed in upscaling/cnn_test.ipynb cells 12
and 18 (tests/test_oracle.py), and the TF "same"-padding / Conv2DTranspose tap alignment, checked against a pure-Python
loop restatement of TensorFlow's documented rules (oracle/tf_loops.py).

Reference lines each function follows (all under /root/reference/upscaling/upscaler/):
  residual_block            model.py:15-27      upsampling_block      model.py:70-75
  make_upscaler_orig        model.py:267-295    VGG_*_LOSS            model.py:101-157
  wasserstein_loss          model.py:159-160    GanLosses family      model.py:166-261
  make_discriminator_simple_512  model.py:836-896
  make_upscaler_attention   model.py:296-328 (residual_block_attention :30-48, upsampling_block_attention :78-98)
  make_upscaler_incep_resnet model.py:445-497 (inception_mini_resblock :372-382, _3path :389-413, _2path :420-443)
  compile helpers / train steps  model.py:1017-1137, train_mini.py:224-234, train_gan.py:296-321, train_gan3.py:337-354
  tensor I/O contract       data.py:253-270

Keras / TF numerics restated here (un-vendored third-party behaviour, Keras 2.2.x + TF 1.14):
  * Conv2D 'same': out = ceil(in/s); pad_total = max((out-1)*s + k - in, 0); pad_before = pad_total // 2.
  * Conv2DTranspose 'same' stride 2: gradient of the conv above, out = 2*in.
  * BatchNormalization: eps 1e-3, momentum 0.99, training normalises with the biased batch variance; moving variance
    is updated with var * n / (n - (1 + eps)).
  * PReLU(shared_axes=[1,2]): alpha of shape (1,1,C).  LeakyReLU alpha 0.2 (generator) / 0.1 (discriminator).
  * Adam(lr=1e-3, b1=0.9, b2=0.999, eps=1e-7), lr_t = lr*sqrt(1-b2^t)/(1-b1^t), p -= lr_t*m/(sqrt(v)+eps).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

BN_EPS = 1e-3
BN_MOMENTUM = 0.99

VGG19_CFG = [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512]
VGG19_NAMES = ["block1_conv1", "block1_conv2", "block2_conv1", "block2_conv2", "block3_conv1", "block3_conv2",
               "block3_conv3", "block3_conv4", "block4_conv1", "block4_conv2", "block4_conv3", "block4_conv4",
               "block5_conv1", "block5_conv2", "block5_conv3", "block5_conv4"]


# ------------------------------------------------------------------------------------------------ primitives
def tf_same_pad(size, k, s):
    """(pad_before, pad_after, out) of TensorFlow's SAME padding."""
    out = -(-size // s)
    total = max((out - 1) * s + k - size, 0)
    return total // 2, total - total // 2, out


def _t(a):
    return a if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a), dtype=torch.float32)


def conv2d_same(x, kernel, bias=None, stride=1):
    """x NCHW; kernel in Keras layout (kh, kw, Cin, Cout)."""
    kh, kw = kernel.shape[0], kernel.shape[1]
    pt, pb, _ = tf_same_pad(x.shape[2], kh, stride)
    pl, pr, _ = tf_same_pad(x.shape[3], kw, stride)
    x = F.pad(x, (pl, pr, pt, pb))
    return F.conv2d(x, kernel.permute(3, 2, 0, 1), bias, stride=stride)


def conv2d_transpose_same_s2(x, kernel, bias=None):
    """x NCHW; kernel in Keras Conv2DTranspose layout (kh, kw, Cout, Cin); stride 2, 'same' => out = 2*in."""
    kh, kw = kernel.shape[0], kernel.shape[1]
    w = kernel.permute(3, 2, 0, 1)  # torch conv_transpose weight: (Cin, Cout, kh, kw)
    y = F.conv_transpose2d(x, w, None, stride=2, padding=0)
    pbh, pbw = (kh - 2) // 2, (kw - 2) // 2
    y = y[:, :, pbh:pbh + 2 * x.shape[2], pbw:pbw + 2 * x.shape[3]]
    if bias is not None:
        y = y + bias.view(1, -1, 1, 1)
    return y


def batch_norm(x, gamma, beta, mean, var, training, stats_out=None):
    """Keras BatchNormalization over the channel axis of an NCHW (or NC) tensor."""
    dims = [0, 2, 3] if x.dim() == 4 else [0]
    shape = (1, -1, 1, 1) if x.dim() == 4 else (1, -1)
    if training:
        m = x.mean(dim=dims)
        v = x.var(dim=dims, unbiased=False)
        if stats_out is not None:
            n = x.numel() // x.shape[1]
            stats_out.append((m.detach(), v.detach(), n))
    else:
        m, v = mean, var
    return (x - m.view(shape)) * (gamma / torch.sqrt(v + BN_EPS)).view(shape) + beta.view(shape)


def update_moving(mean, var, batch_mean, batch_var, n, momentum=BN_MOMENTUM):
    """Keras 2.2.x moving-statistics update (normalization.py: variance *= n / (n - (1 + eps)))."""
    adj = batch_var * (n / (n - (1.0 + BN_EPS)))
    return mean * momentum + batch_mean * (1 - momentum), var * momentum + adj * (1 - momentum)


def prelu(x, alpha):
    a = alpha.reshape(1, -1, 1, 1)
    return torch.clamp(x, min=0) + a * torch.clamp(x, max=0)


def leaky(x, slope):
    return torch.where(x > 0, x, x * slope)


# ------------------------------------------------------------------------------------------------ weights
def glorot_uniform(rng, shape, fan_in, fan_out):
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def _conv_w(rng, kh, kw, cin, cout, bias_scale=0.0):
    k = glorot_uniform(rng, (kh, kw, cin, cout), kh * kw * cin, kh * kw * cout)
    b = (rng.standard_normal(cout) * bias_scale).astype(np.float32)
    return [k, b]


def _bn_w(rng, c, randomize):
    if randomize:
        return [rng.uniform(0.5, 1.5, c).astype(np.float32), (rng.standard_normal(c) * 0.1).astype(np.float32),
                (rng.standard_normal(c) * 0.1).astype(np.float32), rng.uniform(0.5, 1.5, c).astype(np.float32)]
    return [np.ones(c, np.float32), np.zeros(c, np.float32), np.zeros(c, np.float32), np.ones(c, np.float32)]


def init_upscaler_orig_weights(seed=1234, kernel_size=5, filters=64, upscale_factor=4, res_block_num=16, randomize=True):
    """Seeded synthetic weights in Keras layout, keyed by the reference's layer names (model.py:275-291).
    `randomize` perturbs BN statistics, PReLU slopes and biases away from Keras' init so parity is meaningful
    (SURVEY.md 8(d))."""
    rng = np.random.default_rng(seed)
    bs = 0.05 if randomize else 0.0
    w = {}
    w["initial/conv"] = _conv_w(rng, 9, 9, 3, filters, bs)
    w["initial/prelu"] = [(rng.uniform(0, 0.3, (1, 1, filters)) if randomize else np.zeros((1, 1, filters))).astype(np.float32)]
    for i in range(res_block_num):
        n = "res_block/%d" % i
        w[n + "/conv_pre"] = _conv_w(rng, kernel_size, kernel_size, filters, filters, bs)
        w[n + "/batch_norm_pre"] = _bn_w(rng, filters, randomize)
        w[n + "/prelu"] = [(rng.uniform(0, 0.3, (1, 1, filters)) if randomize else np.zeros((1, 1, filters))).astype(np.float32)]
        w[n + "/conv_post"] = _conv_w(rng, kernel_size, kernel_size, filters, filters, bs)
        w[n + "/batch_norm_post"] = _bn_w(rng, filters, randomize)
    w["prefinal/conv2d"] = _conv_w(rng, kernel_size, kernel_size, filters, 64, bs)
    w["prefinal/batch_norm"] = _bn_w(rng, 64, randomize)
    cin = 64
    for i in range(int(math.log(upscale_factor, 2))):
        k = glorot_uniform(rng, (kernel_size, kernel_size, 256, cin), kernel_size * kernel_size * cin,
                           kernel_size * kernel_size * 256)
        w["upscaling/%d/block/conv_transp" % i] = [k, (rng.standard_normal(256) * bs).astype(np.float32)]
        cin = 256
    w["final/conv"] = _conv_w(rng, 9, 9, cin, 3, bs)
    return w


def init_vgg19_weights(seed=4321):
    """He-normal stand-in for the ImageNet weights keras.applications would download (model.py:108)."""
    rng = np.random.default_rng(seed)
    w, cin, i = {}, 3, 0
    for c in VGG19_CFG:
        if c == "M":
            continue
        k = (rng.standard_normal((3, 3, cin, c)) * math.sqrt(2.0 / (9 * cin))).astype(np.float32)
        w[VGG19_NAMES[i]] = [k, (rng.standard_normal(c) * 0.01).astype(np.float32)]
        cin, i = c, i + 1
    return w


DISC_FILTERS = [64, 128, 256, 512, 512, 512, 512, 512, 512]


DISC_THIN_FILTERS = [64, 128, 128, 128, 128, 128, 128, 128, 128]     # make_discriminator_thin_512, model.py:901-939


def init_discriminator_simple_512_weights(seed=99, input_shape=(512, 512, 3), randomize=True, filters=None):
    """`filters`: DISC_FILTERS (simple_512, default) or DISC_THIN_FILTERS (thin_512); the forward pass is the same."""
    rng = np.random.default_rng(seed)
    w, cin = {}, input_shape[2]
    h, wd = input_shape[0], input_shape[1]
    for i, c in enumerate(filters or DISC_FILTERS):
        w["discriminator/block_%d/Conv2d" % (i + 1)] = _conv_w(rng, 3, 3, cin, c, 0.02 if randomize else 0.0)
        w["discriminator/block_%d/BatchNorm" % (i + 1)] = _bn_w(rng, c, randomize)
        if i > 0:
            h, wd = -(-h // 2), -(-wd // 2)
        cin = c
    flat = h * wd * cin
    for j, (fi, fo) in enumerate([(flat, 1024), (1024, 32), (32, 1)]):
        w["discriminator/final/Dense_%d" % (j + 1)] = [glorot_uniform(rng, (fi, fo), fi, fo), np.zeros(fo, np.float32)]
        if j < 2:
            w["discriminator/final/BatchNorm_%d" % (j + 1)] = _bn_w(rng, fo, randomize)
    return w


DISC_SPARSE_FILTERS = [64, 128, 256, 256, 256, 256]       # make_discriminator_sparse_512, model.py:967-989
DISC_SPARSE_STRIDES = [1, 3, 3, 3, 3, 3]                   # 5x5 kernels, padding='valid': 512 -> 508 -> 168 -> 55 -> 17 -> 5 -> 1


def sparse_512_conv_sizes(h, w):
    """Spatial sizes after each 'valid' 5x5 block of make_discriminator_sparse_512 (model.py:967-989)."""
    out = []
    for s in DISC_SPARSE_STRIDES:
        h, w = (h - 5) // s + 1, (w - 5) // s + 1
        if h < 1 or w < 1:
            raise ValueError("input too small for make_discriminator_sparse_512")
        out.append((h, w))
    return out


def init_discriminator_sparse_512_weights(seed=98, input_shape=(512, 512, 3), randomize=True):
    rng = np.random.default_rng(seed)
    w, cin = {}, input_shape[2]
    for i, c in enumerate(DISC_SPARSE_FILTERS):
        w["discriminator/block_%d/Conv2d" % (i + 1)] = _conv_w(rng, 5, 5, cin, c, 0.02 if randomize else 0.0)
        w["discriminator/block_%d/BatchNorm" % (i + 1)] = _bn_w(rng, c, randomize)
        cin = c
    h, wd = sparse_512_conv_sizes(input_shape[0], input_shape[1])[-1]
    flat = h * wd * cin
    for j, (fi, fo) in enumerate([(flat, 128), (128, 32), (32, 1)]):      # model.py:992-1000
        w["discriminator/final/Dense_%d" % (j + 1)] = [glorot_uniform(rng, (fi, fo), fi, fo), np.zeros(fo, np.float32)]
        if j < 2:
            w["discriminator/final/BatchNorm_%d" % (j + 1)] = _bn_w(rng, fo, randomize)
    return w


def to_torch(weights, requires_grad=False):
    out = {}
    for k, arrs in weights.items():
        out[k] = [torch.tensor(np.asarray(a), dtype=torch.float32) for a in arrs]
        if requires_grad:
            for j, t in enumerate(out[k]):
                is_moving = ("atch" in k and ("norm" in k or "Norm" in k)) and j >= 2
                t.requires_grad_(not is_moving)
    return out


def count_params(weights):
    return int(sum(int(np.prod(np.asarray(a).shape)) for arrs in weights.values() for a in arrs))


# ------------------------------------------------------------------------------------------------ generator
def upscaler_orig_forward(w, x_nhwc, training=False, stats_out=None, upscale_factor=4, res_block_num=16):
    """make_upscaler_orig graph (model.py:267-295).  x NHWC in [-1,1]; returns NHWC."""
    x = _t(x_nhwc).permute(0, 3, 1, 2)
    k, b = w["initial/conv"]
    m = conv2d_same(x, k, b)
    m = prelu(m, w["initial/prelu"][0])
    skip = m
    for i in range(res_block_num):
        n = "res_block/%d" % i
        g = m
        m = conv2d_same(m, *w[n + "/conv_pre"])
        m = batch_norm(m, *w[n + "/batch_norm_pre"], training, stats_out)
        m = prelu(m, w[n + "/prelu"][0])
        m = conv2d_same(m, *w[n + "/conv_post"])
        m = batch_norm(m, *w[n + "/batch_norm_post"], training, stats_out)
        m = g + m
    m = conv2d_same(m, *w["prefinal/conv2d"])
    m = batch_norm(m, *w["prefinal/batch_norm"], training, stats_out)
    m = skip + m
    for i in range(int(math.log(upscale_factor, 2))):
        m = conv2d_transpose_same_s2(m, *w["upscaling/%d/block/conv_transp" % i])
        m = leaky(m, 0.2)
    m = conv2d_same(m, *w["final/conv"])
    m = torch.tanh(m)
    return m.permute(0, 2, 3, 1)


def resize_bilinear_tf1(x_nchw, factor):
    """K.resize_images(x, factor, factor, 'channels_last', 'bilinear') as Keras 2.2 / TF 1.14 evaluate it (reference
    model.py:357): tf.image.resize_bilinear with align_corners=False and no half-pixel centres -- source coordinate =
    out * in/out, y0 = floor, y1 = min(y0+1, in-1), lerp in x then y.  (NOT torch's F.interpolate, which uses half-pixel
    centres.)"""
    B, C, H, W = x_nchw.shape
    sy = torch.arange(H * factor, dtype=torch.float32) / factor
    sx = torch.arange(W * factor, dtype=torch.float32) / factor
    y0, x0 = sy.floor().long(), sx.floor().long()
    y1, x1 = torch.clamp(y0 + 1, max=H - 1), torch.clamp(x0 + 1, max=W - 1)
    fy, fx = (sy - y0.float()).view(1, 1, -1, 1), (sx - x0.float()).view(1, 1, 1, -1)
    top = x_nchw[:, :, y0][:, :, :, x0] + (x_nchw[:, :, y0][:, :, :, x1] - x_nchw[:, :, y0][:, :, :, x0]) * fx
    bot = x_nchw[:, :, y1][:, :, :, x0] + (x_nchw[:, :, y1][:, :, :, x1] - x_nchw[:, :, y1][:, :, :, x0]) * fx
    return top + (bot - top) * fy


def init_upscaler_skip_con_weights(seed=1234, kernel_size=5, upscale_factor=4, randomize=True):
    """make_upscaler_skip_con (model.py:332-363) in Keras layout under make_upscaler_orig's layer names."""
    w = init_upscaler_orig_weights(seed=seed, kernel_size=kernel_size, upscale_factor=upscale_factor, randomize=randomize)
    rng = np.random.default_rng(seed + 77)
    w["prefinal/conv2d"] = _conv_w(rng, 3, 3, 64, 64, 0.02 if randomize else 0.0)
    cin = 64
    for i in range(int(math.log(upscale_factor, 2))):
        k = glorot_uniform(rng, (3, 3, 224, cin), 9 * cin, 9 * 224)         # Conv2DTranspose kernel: (kh, kw, Cout, Cin)
        w["upscaling/%d/block/conv_transp" % i] = [k, (rng.normal(0, 0.02, 224) if randomize else np.zeros(224)).astype(np.float32)]
        cin = 224
    w["final/conv"] = _conv_w(rng, 9, 9, 227, 3, 0.02 if randomize else 0.0)
    return w


def upscaler_skip_con_forward(w, x_nhwc, training=False, stats_out=None, upscale_factor=4, res_block_num=16):
    """make_upscaler_skip_con (model.py:332-363); `training` selects batch statistics (with `stats_out` collecting them in
    graph order, as upscaler_orig_forward does) or moving statistics."""
    x = _t(x_nhwc).permute(0, 3, 1, 2)
    m = prelu(conv2d_same(x, *w["initial/conv"]), w["initial/prelu"][0])
    skip = m
    for i in range(res_block_num):
        n = "res_block/%d" % i
        h = conv2d_same(m, *w[n + "/conv_pre"])
        h = prelu(batch_norm(h, *w[n + "/batch_norm_pre"], training, stats_out), w[n + "/prelu"][0])
        h = batch_norm(conv2d_same(h, *w[n + "/conv_post"]), *w[n + "/batch_norm_post"], training, stats_out)
        m = m + h
    m = skip + batch_norm(conv2d_same(m, *w["prefinal/conv2d"]), *w["prefinal/batch_norm"], training, stats_out)
    for i in range(int(math.log(upscale_factor, 2))):
        m = leaky(conv2d_transpose_same_s2(m, *w["upscaling/%d/block/conv_transp" % i]), 0.2)
    m = torch.cat([resize_bilinear_tf1(x, upscale_factor), m], dim=1)      # Concatenate(axis=3) in NHWC
    return torch.tanh(conv2d_same(m, *w["final/conv"])).permute(0, 2, 3, 1)


def conv2d_transpose_same(x, kernel, bias=None, stride=2):
    """Conv2DTranspose(padding='same') for any stride <= kernel: the gradient of the SAME conv that maps stride*in -> in
    (pad_total = kernel - stride, pad_before = pad_total // 2), out = stride * in.  x NCHW; kernel (kh, kw, Cout, Cin)."""
    kh, kw = kernel.shape[0], kernel.shape[1]
    y = F.conv_transpose2d(x, kernel.permute(3, 2, 0, 1), None, stride=stride, padding=0)
    pbh, pbw = max(kh - stride, 0) // 2, max(kw - stride, 0) // 2
    y = y[:, :, pbh:pbh + stride * x.shape[2], pbw:pbw + stride * x.shape[3]]
    if bias is not None:
        y = y + bias.view(1, -1, 1, 1)
    return y


def resize_nearest_tf1(x_nchw, factor):
    """K.resize_images(..., 'nearest') = tf.image.resize_nearest_neighbor, align_corners=False: source = out // factor."""
    return x_nchw.repeat_interleave(factor, dim=2).repeat_interleave(factor, dim=3)


ATT_UP_FILTERS = 128      # make_upscaler_attention passes filters=128 to its upsampling blocks (model.py:319)


def init_upscaler_attention_weights(seed=1234, kernel_size=5, filters=64, upscale_factor=4, res_block_num=16, randomize=True):
    """make_upscaler_attention (model.py:296-328) in Keras layout under the reference's layer names."""
    rng = np.random.default_rng(seed)
    bs = 0.05 if randomize else 0.0
    k = kernel_size
    alpha = lambda: [(rng.uniform(0, 0.3, (1, 1, filters)) if randomize else np.zeros((1, 1, filters))).astype(np.float32)]
    w = {"initial/conv": _conv_w(rng, 9, 9, 3, filters, bs), "initial/prelu": alpha()}
    for i in range(res_block_num):
        n = "res_block/%d" % i
        w[n + "/attention"] = _conv_w(rng, k, k, 3, filters, bs)
        w[n + "/conv_pre"] = _conv_w(rng, k, k, filters, filters, bs)
        w[n + "/batch_norm_pre"] = _bn_w(rng, filters, randomize)
        w[n + "/prelu"] = alpha()
        w[n + "/conv_post"] = _conv_w(rng, k, k, filters, filters, bs)
        w[n + "/batch_norm_post"] = _bn_w(rng, filters, randomize)
    w["after_res/conv"] = _conv_w(rng, k, k, filters, filters, bs)
    w["after_res/batch_norm"] = _bn_w(rng, filters, randomize)
    cin = filters
    for i in range(int(math.log(upscale_factor, 2))):
        n, scale = "upscaling/%d/block" % i, 2 ** (i + 1)
        w[n + "/attention"] = _conv_w(rng, k, k, 6, cin, bs)
        w[n + "/conv_transp"] = [glorot_uniform(rng, (k, k, ATT_UP_FILTERS, cin), k * k * cin, k * k * ATT_UP_FILTERS),
                                 (rng.standard_normal(ATT_UP_FILTERS) * bs).astype(np.float32)]
        ks = scale + 1
        w[n + "/to_add_input_conv_transp"] = [glorot_uniform(rng, (ks, ks, ATT_UP_FILTERS, 3), ks * ks * 3, ks * ks * ATT_UP_FILTERS),
                                              (rng.standard_normal(ATT_UP_FILTERS) * bs).astype(np.float32)]
        cin = ATT_UP_FILTERS
    w["final/conv"] = _conv_w(rng, 9, 9, cin, 3, bs)
    return w


def upscaler_attention_forward(w, x_nhwc, training=False, stats_out=None, upscale_factor=4, res_block_num=16):
    """make_upscaler_attention (model.py:296-328) with residual_block_attention (:30-48) and upsampling_block_attention
    (:78-98): every residual block's input is gated by sigmoid(conv(network input)); each upsampling block gates its
    input by sigmoid(conv(Concatenate([nearest, bilinear] resize of the network input by scale // 2))), applies
    Conv2DTranspose(128, k, stride 2) + LeakyReLU(0.2) and adds Conv2DTranspose(128, scale + 1, stride scale) of
    atanh(0.99999 * network input)."""
    x = _t(x_nhwc).permute(0, 3, 1, 2)
    m = prelu(conv2d_same(x, *w["initial/conv"]), w["initial/prelu"][0])
    skip = m
    for i in range(res_block_num):
        n = "res_block/%d" % i
        h = torch.sigmoid(conv2d_same(x, *w[n + "/attention"])) * m
        h = conv2d_same(h, *w[n + "/conv_pre"])
        h = prelu(batch_norm(h, *w[n + "/batch_norm_pre"], training, stats_out), w[n + "/prelu"][0])
        h = batch_norm(conv2d_same(h, *w[n + "/conv_post"]), *w[n + "/batch_norm_post"], training, stats_out)
        m = m + h
    m = skip + batch_norm(conv2d_same(m, *w["after_res/conv"]), *w["after_res/batch_norm"], training, stats_out)
    for i in range(int(math.log(upscale_factor, 2))):
        n, scale = "upscaling/%d/block" % i, 2 ** (i + 1)
        r = scale // 2
        iu = torch.cat([resize_nearest_tf1(x, r), resize_bilinear_tf1(x, r)], dim=1)
        m = torch.sigmoid(conv2d_same(iu, *w[n + "/attention"])) * m
        m = leaky(conv2d_transpose_same(m, *w[n + "/conv_transp"], stride=2), 0.2)
        m = m + conv2d_transpose_same(torch.atanh(0.99999 * x), *w[n + "/to_add_input_conv_transp"], stride=scale)
    return torch.tanh(conv2d_same(m, *w["final/conv"])).permute(0, 2, 3, 1)


def upscaler_attention_bn_names(res_block_num=16):
    names = []
    for i in range(res_block_num):
        names += ["res_block/%d/batch_norm_pre" % i, "res_block/%d/batch_norm_post" % i]
    return names + ["after_res/batch_norm"]


def incep_resnet_blocks(filters=64, a=("3path", 5, 3), b=("2path", 10, 7), c=("2path", 5, 3)):
    """[(block name, [(path, [(mini-block name, (kh, kw), cin, cout), ...]), ...], concat width)] in graph order for
    make_upscaler_incep_resnet (model.py:389-443, 445-497): the three groups are named inc_res_block/A, /B and /c."""
    f = filters
    out = []
    for tag, (btype, num, k) in (("A", a), ("B", b), ("c", c)):
        for i in range(num):
            if btype == "3path":
                base = "inc_res_block/%s/3p/%d" % (tag, i)
                f1, f2, f3 = int(f * 0.5), int(f * 0.75), f
                paths = [("a", [(base + "/a/1", (1, 1), f, f1)]),
                         ("b", [(base + "/b/1", (1, 1), f, f1), (base + "/b/2", (k, k), f1, f1)]),
                         ("c", [(base + "/c/1", (1, 1), f, f1), (base + "/c/2", (k, k), f1, f2), (base + "/c/3", (k, k), f2, f3)])]
            else:
                base = "inc_res_block/%s/2p/%d" % (tag, i)
                fa, f1, f2, f3 = int(f * 0.5), int(f * 0.3), int(f * 0.4), int(f * 0.5)
                paths = [("a", [(base + "/a/1", (1, 1), f, fa)]),
                         ("b", [(base + "/b/1", (1, 1), f, f1), (base + "/b/2", (1, k), f1, f2), (base + "/b/3", (k, 1), f2, f3)])]
            out.append((base, paths, sum(p[1][-1][3] for p in paths)))
    return out


def init_upscaler_incep_resnet_weights(seed=1234, filters=64, upscale_factor=4, a=("3path", 5, 3), b=("2path", 10, 7),
                                       c=("2path", 5, 3), randomize=True, block_gain=0.5):
    """make_upscaler_incep_resnet (model.py:445-497) in Keras layout under the reference's layer names: every mini block is
    BatchNormalization -> PReLU -> Conv2D named <name>/batch_norm, <name>/prelu, <name>/<kh>x<kw> (model.py:372-382)."""
    rng = np.random.default_rng(seed)
    bs = 0.05 if randomize else 0.0
    f = filters
    w = {"initial/conv/9x9": _conv_w(rng, 9, 9, 3, f, bs)}
    for base, paths, cat in incep_resnet_blocks(f, a, b, c):
        for _, minis in paths:
            for name, (kh, kw), cin, cout in minis:
                w[name + "/batch_norm"] = _bn_w(rng, cin, randomize)
                w[name + "/prelu"] = [(rng.uniform(0, 0.3, (1, 1, cin)) if randomize else np.zeros((1, 1, cin))).astype(np.float32)]
                w[name + "/%dx%d" % (kh, kw)] = _conv_w(rng, kh, kw, cin, cout, bs)
        # block_gain scales the 1x1 conv that closes a block: with glorot kernels at gain 1 the 20 pre-activation blocks
        # push the residual stream to std ~20 and 13 % of the outputs into tanh saturation, where nothing can be compared
        w[base + "/final/1x1"] = _conv_w(rng, 1, 1, cat, f, bs)
        w[base + "/final/1x1"][0] = (w[base + "/final/1x1"][0] * block_gain).astype(np.float32)
    ck = c[2]
    w["prefinal/conv2d"] = _conv_w(rng, ck, ck, f, f, bs)
    w["prefinal/batch_norm"] = _bn_w(rng, f, randomize)
    cin = f
    for i in range(int(math.log(upscale_factor, 2))):
        w["upscaling/%d/block/conv_transp" % i] = [glorot_uniform(rng, (ck, ck, 256, cin), ck * ck * cin, ck * ck * 256),
                                                   (rng.standard_normal(256) * bs).astype(np.float32)]
        cin = 256
    w["final/conv"] = _conv_w(rng, 9, 9, cin, 3, bs)
    return w


def upscaler_incep_resnet_forward(w, x_nhwc, training=False, stats_out=None, filters=64, upscale_factor=4, a=("3path", 5, 3),
                                  b=("2path", 10, 7), c=("2path", 5, 3)):
    """make_upscaler_incep_resnet (model.py:445-497): 9x9 conv (no activation) -> inception residual blocks (each path a chain
    of BN -> PReLU -> conv mini blocks, the paths concatenated, a 1x1 conv back to `filters`, added to the block input) ->
    conv + BN + skip -> two Conv2DTranspose(256, k, stride 2) + LeakyReLU(0.2) -> 9x9 conv -> tanh."""
    x = _t(x_nhwc).permute(0, 3, 1, 2)
    m = conv2d_same(x, *w["initial/conv/9x9"])
    skip = m
    for base, paths, _ in incep_resnet_blocks(filters, a, b, c):
        outs = []
        for _, minis in paths:
            t = m
            for name, (kh, kw), _, _ in minis:
                t = prelu(batch_norm(t, *w[name + "/batch_norm"], training, stats_out), w[name + "/prelu"][0])
                t = conv2d_same(t, *w[name + "/%dx%d" % (kh, kw)])
            outs.append(t)
        m = m + conv2d_same(torch.cat(outs, dim=1), *w[base + "/final/1x1"])
    m = skip + batch_norm(conv2d_same(m, *w["prefinal/conv2d"]), *w["prefinal/batch_norm"], training, stats_out)
    for i in range(int(math.log(upscale_factor, 2))):
        m = leaky(conv2d_transpose_same_s2(m, *w["upscaling/%d/block/conv_transp" % i]), 0.2)
    return torch.tanh(conv2d_same(m, *w["final/conv"])).permute(0, 2, 3, 1)


def upscaler_bn_names(res_block_num=16):
    names = []
    for i in range(res_block_num):
        names += ["res_block/%d/batch_norm_pre" % i, "res_block/%d/batch_norm_post" % i]
    return names + ["prefinal/batch_norm"]


# ------------------------------------------------------------------------------------------------ VGG19 + losses
def bf16_ste(x):
    """Round to bf16 in the forward pass, identity in the backward pass (storage emulation for gradient tests)."""
    return x + (x.to(torch.bfloat16).to(torch.float32) - x).detach()


def f16_ste(x):
    """Round to IEEE fp16 in the forward pass, identity in the backward pass."""
    return x + (x.to(torch.float16).to(torch.float32) - x).detach()


def vgg19_block5_conv4(vw, x_nhwc, q=None):
    """keras.applications VGG19(include_top=False) up to block5_conv4 (post-ReLU); raw [-1,1] input (model.py:117).
    `q` (tests only) is applied to every stored activation, e.g. bf16_ste to emulate 16-bit activation storage so that
    ReLU gates are decided on the same rounded values as in the CUDA path."""
    x = _t(x_nhwc).permute(0, 3, 1, 2)
    if q is not None:
        x = q(x)
    i = 0
    for c in VGG19_CFG:
        if c == "M":
            x = F.max_pool2d(x, 2, 2)
        else:
            k, b = vw[VGG19_NAMES[i]]
            x = F.relu(conv2d_same(x, q(k) if q is not None else k, b))
            if q is not None:
                x = q(x)
            i += 1
    return x.permute(0, 2, 3, 1)


def content_loss(kind, vw, y_true, y_pred, rate):
    """kind in {'vgg', 'vgg-mse', 'vgg-mae', 'mse', 'mae'} (model.py:101-157; train.py -l choices)."""
    y_true, y_pred = _t(y_true), _t(y_pred)
    if kind == "mse":
        return torch.mean((y_true - y_pred) ** 2)
    if kind == "mae":
        return torch.mean(torch.abs(y_true - y_pred))
    ft, fp = vgg19_block5_conv4(vw, y_true), vgg19_block5_conv4(vw, y_pred)
    if kind == "vgg":
        return torch.mean((ft - fp) ** 2)
    if kind == "vgg-mse":
        return torch.mean((ft - fp) ** 2) + rate * torch.mean((y_true - y_pred) ** 2)
    if kind == "vgg-mae":
        return torch.mean(torch.abs(ft - fp)) + rate * torch.mean(torch.abs(y_true - y_pred))
    raise ValueError(kind)


def wasserstein_loss(y_true, y_pred):
    return torch.mean(_t(y_true) * _t(y_pred))


def loss_activation(name, x):
    """GanLosses.__init__ activation table (model.py:170-181); anything else is identity."""
    if name == "sigmoid":
        return torch.sigmoid(x)
    if name == "log-sigm":
        return torch.log(torch.sigmoid(x))
    if name == "tanh":
        return torch.tanh(x)
    if name == "bi-log":
        return (x / (1 + torch.abs(x))) * torch.log(torch.abs(x) + 2)
    return x


def gan_losses(family, act, real_out, fake_out):
    """(discriminator_loss, generator_loss) of WassersteinLosses / RelativisticLosses (model.py:215-261)."""
    if family == "wasserstein":
        return torch.mean(real_out) - torch.mean(fake_out), torch.mean(fake_out)
    d = loss_activation(act, torch.mean(real_out) - torch.mean(fake_out))
    g = loss_activation(act, torch.mean(fake_out) - torch.mean(real_out))
    return d, g


# ------------------------------------------------------------------------------------------------ discriminator
def discriminator_simple_512_forward(w, x_nhwc, activation="none", training=False, stats_out=None, q=None):
    """make_discriminator_simple_512 (model.py:836-896) and make_discriminator_thin_512 (model.py:901-960: same graph,
    narrower blocks -- the widths come with the weights); returns (n, 1).  `q` (tests only): applied to every tensor the
    CUDA path stores in 16 bits (conv / dense outputs and activations, conv weights), e.g. f16_ste -- the fp32 oracle's own
    sensitivity to storage rounding, from which the adversarial-loss tolerances of the GAN tests are derived."""
    qq = q if q is not None else (lambda t: t)
    x = qq(_t(x_nhwc).permute(0, 3, 1, 2))
    for i in range(9):
        k, b = w["discriminator/block_%d/Conv2d" % (i + 1)]
        x = qq(conv2d_same(x, qq(k), b, stride=1 if i == 0 else 2))
        x = batch_norm(x, *w["discriminator/block_%d/BatchNorm" % (i + 1)], training, stats_out)
        x = qq(leaky(x, 0.1))
    x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)  # Keras Flatten of an NHWC tensor
    for j in range(2):
        k, b = w["discriminator/final/Dense_%d" % (j + 1)]
        x = qq(x @ k + b)
        x = batch_norm(x, *w["discriminator/final/BatchNorm_%d" % (j + 1)], training, stats_out)
        x = qq(leaky(x, 0.1))
    k, b = w["discriminator/final/Dense_3"]
    x = x @ k + b
    return loss_activation(activation, x) if activation in ("sigmoid", "log-sigm", "tanh", "bi-log") else x


def discriminator_sparse_512_forward(w, x_nhwc, activation="none", training=False, stats_out=None):
    """make_discriminator_sparse_512 (model.py:964-1012): six 5x5 padding='valid' blocks (strides 1,3,3,3,3,3) + the
    Dense 128 / 32 / 1 head; returns (n, 1)."""
    x = _t(x_nhwc).permute(0, 3, 1, 2)
    for i, s in enumerate(DISC_SPARSE_STRIDES):
        k, b = w["discriminator/block_%d/Conv2d" % (i + 1)]
        x = F.conv2d(x, k.permute(3, 2, 0, 1), b, stride=s)
        x = batch_norm(x, *w["discriminator/block_%d/BatchNorm" % (i + 1)], training, stats_out)
        x = leaky(x, 0.1)
    x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
    for j in range(2):
        k, b = w["discriminator/final/Dense_%d" % (j + 1)]
        x = x @ k + b
        x = batch_norm(x, *w["discriminator/final/BatchNorm_%d" % (j + 1)], training, stats_out)
        x = leaky(x, 0.1)
    k, b = w["discriminator/final/Dense_3"]
    x = x @ k + b
    return loss_activation(activation, x) if activation in ("sigmoid", "log-sigm", "tanh", "bi-log") else x


def discriminator_bn_names(weights=None):
    """BatchNormalization layers in graph order; the block count (9, or 6 for sparse_512) is read from `weights`."""
    nb = 9 if weights is None else len([k for k in weights if k.startswith("discriminator/block_") and k.endswith("/BatchNorm")])
    return ["discriminator/block_%d/BatchNorm" % (i + 1) for i in range(nb)] + \
           ["discriminator/final/BatchNorm_1", "discriminator/final/BatchNorm_2"]


# ------------------------------------------------------------------------------------------------ Keras Adam
class KerasAdam:
    """keras.optimizers.Adam defaults; `iterations` may be shared between several trainers."""

    def __init__(self, lr=1e-3, beta_1=0.9, beta_2=0.999, epsilon=1e-7):
        self.lr, self.b1, self.b2, self.eps = lr, beta_1, beta_2, epsilon
        self.iterations = 0
        self.slots = {}

    def apply(self, params):
        """params: list of (key, tensor with .grad). One call == one Keras update op."""
        self.iterations += 1
        t = self.iterations
        lr_t = self.lr * math.sqrt(1.0 - self.b2 ** t) / (1.0 - self.b1 ** t)
        with torch.no_grad():
            for key, p in params:
                if p.grad is None:
                    continue
                if key not in self.slots:
                    self.slots[key] = (torch.zeros_like(p), torch.zeros_like(p))
                m, v = self.slots[key]
                m.mul_(self.b1).add_(p.grad, alpha=1 - self.b1)
                v.mul_(self.b2).addcmul_(p.grad, p.grad, value=1 - self.b2)
                p.sub_(lr_t * m / (torch.sqrt(v) + self.eps))
                p.grad = None


def _trainable(w, prefix):
    out = []
    for k, ts in w.items():
        for j, t in enumerate(ts):
            if t.requires_grad:
                out.append(("%s:%s:%d" % (prefix, k, j), t))
    return out


def _apply_moving(w, names, stats):
    with torch.no_grad():
        for n, (bm, bv, cnt) in zip(names, stats):
            mm, mv = update_moving(w[n][2], w[n][3], bm, bv, cnt)
            w[n][2].copy_(mm)
            w[n][3].copy_(mv)


class GeneratorTrainer:
    """compile_training_model + train_on_batch (model.py:1130-1137; train_mini.py:232)."""

    def __init__(self, gen_weights, vgg_weights, loss_kind="vgg-mse", rate=1e-3, optimizer=None, res_block_num=16,
                 upscale_factor=4, forward=None, grad_noise=None):
        self.forward = forward or upscaler_orig_forward        # or upscaler_skip_con_forward (same signature)
        # grad_noise = (relative L2 size, seed): every gradient tensor gets white noise of that relative norm before the
        # update.  Tests use it to measure how far the oracle's OWN trajectory moves under a perturbation of the size 16-bit
        # storage causes in the CUDA path (the spread a free-running comparison has to be judged against).
        self.grad_noise = grad_noise
        self._noise_gen = torch.Generator().manual_seed(int(grad_noise[1])) if grad_noise else None
        self.w = to_torch(gen_weights, requires_grad=True)
        self.vw = to_torch(vgg_weights) if vgg_weights is not None else None
        self.kind, self.rate = loss_kind, rate
        self.opt = optimizer or KerasAdam()
        self.rbn, self.uf = res_block_num, upscale_factor

    def train_on_batch(self, x_lr, y_hr):
        stats = []
        y = self.forward(self.w, x_lr, True, stats, self.uf, self.rbn)
        loss = content_loss(self.kind, self.vw, _t(y_hr), y, self.rate)
        loss.backward()
        if self.grad_noise:
            with torch.no_grad():
                for _, p in _trainable(self.w, "g"):
                    if p.grad is not None and p.grad.numel() > 0:
                        n = torch.randn(p.grad.shape, generator=self._noise_gen)
                        p.grad += n * (self.grad_noise[0] * p.grad.norm() / max(float(n.norm()), 1e-30))
        self.opt.apply(_trainable(self.w, "g"))
        _apply_moving(self.w, upscaler_bn_names(self.rbn), stats)
        return float(loss.detach())

    def predict(self, x_lr):
        with torch.no_grad():
            return self.forward(self.w, x_lr, False, None, self.uf, self.rbn).numpy()


class GanTrainer:
    """make_and_compile_gan + the train_gan.py loop body (model.py:1017-1051; train_gan.py:305-317).
    One Adam instance is shared by the discriminator and GAN trainers (default-argument object) so `iterations`
    advances on both calls."""

    def __init__(self, gen_weights, disc_weights, vgg_weights, content_kind="vgg-mse", rate=1e-3, content_weight=1.0,
                 disc_loss_weight=1e-10, disc_activation="none", share_adam_counter=True, res_block_num=16,
                 upscale_factor=4, disc_forward=None, frozen_updates_moving=False):
        # frozen_updates_moving: whether the discriminator's BatchNorm moving statistics follow the training-phase passes.
        # UNPINNED Keras behaviour (no Keras here): from 2.1.3 on, Layer.updates / Network.updates are [] while `trainable` is
        # False, the train functions are built at the first train_on_batch, and make_and_compile_gan* leave
        # discriminator.trainable = False (model.py:1040, 1101) -> False is the default reading; True = pre-2.1.3.
        self.d_moves = frozen_updates_moving
        self.dfwd = disc_forward or discriminator_simple_512_forward     # or discriminator_sparse_512_forward
        self.g = to_torch(gen_weights, requires_grad=True)
        self.d = to_torch(disc_weights, requires_grad=True)
        self.vw = to_torch(vgg_weights) if vgg_weights is not None else None
        self.kind, self.rate, self.cw, self.dw = content_kind, rate, content_weight, disc_loss_weight
        self.dact = disc_activation
        self.opt_d = KerasAdam()
        self.opt_g = self.opt_d if share_adam_counter else KerasAdam()
        self.rbn, self.uf = res_block_num, upscale_factor

    def predict(self, x_lr):
        with torch.no_grad():
            return upscaler_orig_forward(self.g, x_lr, False, None, self.uf, self.rbn).numpy()

    def disc_train_on_batch(self, d_in, d_y):
        stats = []
        out = self.dfwd(self.d, d_in, self.dact, True, stats)
        loss = wasserstein_loss(_t(d_y).reshape(-1, 1), out)
        loss.backward()
        self.last_d_grads = {k: [None if t.grad is None else t.grad.detach().clone() for t in ts] for k, ts in self.d.items()}
        self.opt_d.apply(_trainable(self.d, "d"))
        if self.d_moves:
            _apply_moving(self.d, discriminator_bn_names(self.d), stats)
        return float(loss.detach())

    def gan_train_on_batch(self, x_lr, y_hr, adv_y):
        gstats, dstats = [], []
        y = upscaler_orig_forward(self.g, x_lr, True, gstats, self.uf, self.rbn)
        dout = self.dfwd(self.d, y, self.dact, True, dstats)
        lc = content_loss(self.kind, self.vw, _t(y_hr), y, self.rate)
        la = wasserstein_loss(_t(adv_y).reshape(-1, 1), dout)
        total = self.cw * lc + self.dw * la
        total.backward()
        for ts in self.d.values():  # discriminator frozen inside gan_train (model.py:1040)
            for t in ts:
                t.grad = None
        self.opt_g.apply(_trainable(self.g, "g"))
        _apply_moving(self.g, upscaler_bn_names(self.rbn), gstats)
        # a frozen model's BatchNormalization still normalises with batch statistics (learning phase 1)
        if self.d_moves:
            _apply_moving(self.d, discriminator_bn_names(self.d), dstats)
        return [float(total.detach()), float(lc.detach()), float(la.detach())]

    def step(self, x_lr, y_hr):
        """train_gan.py:305-317 with labels +1 real / -1 fake."""
        n = x_lr.shape[0]
        g = self.predict(x_lr)
        d_in = np.concatenate([np.asarray(y_hr, np.float32), g], axis=0)
        d_y = np.concatenate([np.ones(n, np.float32), -np.ones(n, np.float32)])
        ld = self.disc_train_on_batch(d_in, d_y)
        lg = self.gan_train_on_batch(x_lr, y_hr, np.ones(n, np.float32))
        return ld, lg


class Gan2Trainer(GanTrainer):
    """make_and_compile_gan2 + the train_gan3.py loop body (model.py:1057-1125; train_gan3.py:339-354): the
    discriminator model takes [real, fake] as two inputs (two separate BatchNorm batches) and its loss is
    GanLosses.discriminator_loss of the two symbolic outputs; the GAN model takes [x_lr, real] and adds
    GanLosses.generator_loss, the discriminator being frozen."""

    def __init__(self, *args, family="relativistic", loss_act="log-sigm", **kw):
        super().__init__(*args, **kw)
        self.family, self.loss_act = family, loss_act

    def disc_train_on_batch2(self, real, fake):
        s1, s2 = [], []
        r = self.dfwd(self.d, real, self.dact, True, s1)
        f = self.dfwd(self.d, fake, self.dact, True, s2)
        loss, _ = gan_losses(self.family, self.loss_act, r, f)
        loss.backward()
        self.last_d_grads = {k: [None if t.grad is None else t.grad.detach().clone() for t in ts] for k, ts in self.d.items()}
        self.opt_d.apply(_trainable(self.d, "d"))
        if self.d_moves:
            _apply_moving(self.d, discriminator_bn_names(self.d), s1)
            _apply_moving(self.d, discriminator_bn_names(self.d), s2)
        return float(loss.detach())

    def gan_train_on_batch2(self, x_lr, real, y_hr):
        gstats, s1, s2 = [], [], []
        r = self.dfwd(self.d, real, self.dact, True, s1)
        y = upscaler_orig_forward(self.g, x_lr, True, gstats, self.uf, self.rbn)
        f = self.dfwd(self.d, y, self.dact, True, s2)
        lc = content_loss(self.kind, self.vw, _t(y_hr), y, self.rate)
        _, la = gan_losses(self.family, self.loss_act, r, f)
        total = self.cw * lc + self.dw * la
        total.backward()
        for ts in self.d.values():
            for t in ts:
                t.grad = None
        self.opt_g.apply(_trainable(self.g, "g"))
        _apply_moving(self.g, upscaler_bn_names(self.rbn), gstats)
        if self.d_moves:
            _apply_moving(self.d, discriminator_bn_names(self.d), s1)
            _apply_moving(self.d, discriminator_bn_names(self.d), s2)
        return [float(total.detach()), float(lc.detach()), float(la.detach())]


# ------------------------------------------------------------------------------------------------ tensor I/O contract
def convert_image_to_array(img_u8):
    """data.py:259-263."""
    return np.asarray(img_u8) / 127.5 - 1


def convert_array_to_u8(array):
    """data.py:253-256 without the PIL wrapper."""
    return np.uint8(np.around((np.asarray(array) + 1) * 127.5))


# ------------------------------------------------------------------------------------------------ bf16 emulation
def bf16_round(t):
    return t.to(torch.bfloat16).to(torch.float32)
