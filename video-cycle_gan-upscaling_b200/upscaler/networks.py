"""Device-side networks of the hot path, executed layer by layer through the C ABI.

UpscalerOrigNet is the graph make_upscaler_orig declares (reference: upscaling/upscaler/model.py:267-295, with
residual_block :15-27 and upsampling_block :70-75); Vgg19Net the frozen feature extractor behind the VGG losses
(model.py:101-157); DiscriminatorSimple512Net the graph of make_discriminator_simple_512 (model.py:836-896).

Memory layout: all parameters of a network live in ONE flat fp32 buffer (trainable tensors first, BatchNorm moving
statistics last) with per-layer views in Keras layout under the reference's layer names, so that .h5 groups map
one-to-one, the Adam update is a single launch over the flat buffer, and the data-parallel gradient all-reduce runs
over contiguous slices of the matching flat gradient buffer.  The 16-bit packed copies the tensor-core kernels read
(forward layouts, and flipped / transposed layouts for the data gradients) are refreshed from the masters by one
gather launch per element type.
"""
import math
from collections import OrderedDict

import numpy as np
import torch

from . import ops

# 16-bit storage type of the TRAINING passes of the generator and the discriminators (activations, packed weights and
# gradients; every kernel takes the element type per call).  IEEE fp16 keeps 11 significand bits against bf16's 8: measured
# on B200 against the fp32 oracle (profiles/r02_s4_train_numerics.log, 16 residual blocks) the generator's weight gradients
# sit at a median relative L2 error of 0.14 (worst cosine 0.978) instead of 0.35 (0.872), the first loss at 6e-5 instead of
# 9e-4, and the direction of the two-pass critic gradient g(real) - g(fake) survives (cosine 0.99 instead of 0.3-0.6, which
# is what the fp32 oracle itself gives once its activations are rounded to bf16: profiles/r02_s4_disc2_grad_check.log).
# fp16's narrow range is handled on both sides: gradients are carried multiplied by a per-step scale (training.GRAD_SCALE),
# and the forward range is watched (RANGE_LIMIT below): if it is ever approached the network drops to bf16 storage.
TRAIN_DTYPE = torch.float16
# fp16 training storage stores with a saturating conversion (+-65504).  Every train-mode BatchNorm finalize raises a device
# flag when a channel of its input has a root mean square above RANGE_LIMIT (1/32 of the edge), and the step samples the
# tensors no BatchNorm sees (head conv, Conv2DTranspose outputs, generated image) with count_saturated every
# RANGE_CHECK_EVERY steps; the backward passes count saturated values in the (small) activation-gradient tensors every step.
# Flag and gradient count are the network's two "guard words": the step's Adam launch reads them on the device and drops an
# update computed from clipped values, and the host reads them with every step's loss and switches that network to bf16
# storage with a warning (ON_RANGE = "fallback") or raises FloatingPointError ("raise").
RANGE_LIMIT = 2048.0
RANGE_CHECK_EVERY = 32
GRAD_PROBE_MAX_ELEMS = 1 << 25     # activation gradients up to this size are sampled for saturation every step (<= 64 MB read)
ON_RANGE = "fallback"

BN_EPS = 1e-3       # keras BatchNormalization defaults, model.py:20
BN_MOMENTUM = 0.99
ALIGN = 64          # floats; keeps every tensor (and its gradient) 256-byte aligned inside the flat buffers


def glorot_uniform(rng, shape, fan_in, fan_out):
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def upscaler_orig_layer_specs(kernel_size=5, filters=64, upscale_factor=4, res_block_num=16, up_filters=256,
                              up_kernel=None, prefinal_kernel=None, concat_input=False):
    """(name, kind, config) in graph order, with the reference's layer names.  The keyword variants describe
    make_upscaler_skip_con (model.py:332-363): 3x3 after-res conv, 3x3 Conv2DTranspose with 224 filters, and the final
    conv reading Concatenate([bilinear x4 input (3), upsampled features (224)])."""
    k = kernel_size
    uk, pk_ = up_kernel or k, prefinal_kernel or k
    specs = [("initial/conv", "conv", dict(kh=9, kw=9, cin=3, cout=filters)),
             ("initial/prelu", "prelu", dict(c=filters))]
    for i in range(res_block_num):
        n = "res_block/%d" % i
        specs += [(n + "/conv_pre", "conv", dict(kh=k, kw=k, cin=filters, cout=filters)),
                  (n + "/batch_norm_pre", "bn", dict(c=filters)),
                  (n + "/prelu", "prelu", dict(c=filters)),
                  (n + "/conv_post", "conv", dict(kh=k, kw=k, cin=filters, cout=filters)),
                  (n + "/batch_norm_post", "bn", dict(c=filters))]
    specs += [("prefinal/conv2d", "conv", dict(kh=pk_, kw=pk_, cin=filters, cout=64)),
              ("prefinal/batch_norm", "bn", dict(c=64))]
    cin = 64
    for i in range(int(math.log(upscale_factor, 2))):
        specs.append(("upscaling/%d/block/conv_transp" % i, "convT", dict(kh=uk, kw=uk, cin=cin, cout=up_filters)))
        cin = up_filters
    specs.append(("final/conv", "conv", dict(kh=9, kw=9, cin=cin + (3 if concat_input else 0), cout=3)))
    return specs


ATT_UP_FILTERS = 128      # make_upscaler_attention passes filters = 128 to its upsampling blocks (model.py:319)


def upscaler_attention_layer_specs(kernel_size=5, filters=64, upscale_factor=4, res_block_num=16):
    """make_upscaler_attention (model.py:296-328, blocks :30-48 and :78-98) in graph order with the reference's layer names."""
    k = kernel_size
    specs = [("initial/conv", "conv", dict(kh=9, kw=9, cin=3, cout=filters)),
             ("initial/prelu", "prelu", dict(c=filters))]
    for i in range(res_block_num):
        n = "res_block/%d" % i
        specs += [(n + "/attention", "conv", dict(kh=k, kw=k, cin=3, cout=filters)),
                  (n + "/conv_pre", "conv", dict(kh=k, kw=k, cin=filters, cout=filters)),
                  (n + "/batch_norm_pre", "bn", dict(c=filters)),
                  (n + "/prelu", "prelu", dict(c=filters)),
                  (n + "/conv_post", "conv", dict(kh=k, kw=k, cin=filters, cout=filters)),
                  (n + "/batch_norm_post", "bn", dict(c=filters))]
    specs += [("after_res/conv", "conv", dict(kh=k, kw=k, cin=filters, cout=filters)),
              ("after_res/batch_norm", "bn", dict(c=filters))]
    cin = filters
    for i in range(int(math.log(upscale_factor, 2))):
        n, scale = "upscaling/%d/block" % i, 2 ** (i + 1)
        specs += [(n + "/attention", "conv", dict(kh=k, kw=k, cin=6, cout=cin)),
                  (n + "/conv_transp", "convT", dict(kh=k, kw=k, cin=cin, cout=ATT_UP_FILTERS)),
                  (n + "/to_add_input_conv_transp", "convT", dict(kh=scale + 1, kw=scale + 1, cin=3, cout=ATT_UP_FILTERS))]
        cin = ATT_UP_FILTERS
    specs.append(("final/conv", "conv", dict(kh=9, kw=9, cin=cin, cout=3)))
    return specs


def incep_resnet_blocks(filters=64, a=("3path", 5, 3), b=("2path", 10, 7), c=("2path", 5, 3)):
    """[(block name, [(path, [(mini-block name, (kh, kw), cin, cout), ...]), ...], concat width)] in graph order for
    make_upscaler_incep_resnet (model.py:389-443, 445-497); the three groups are named inc_res_block/A, /B and /c."""
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
            elif btype == "2path":
                base = "inc_res_block/%s/2p/%d" % (tag, i)
                fa, f1, f2, f3 = int(f * 0.5), int(f * 0.3), int(f * 0.4), int(f * 0.5)
                paths = [("a", [(base + "/a/1", (1, 1), f, fa)]),
                         ("b", [(base + "/b/1", (1, 1), f, f1), (base + "/b/2", (1, k), f1, f2), (base + "/b/3", (k, 1), f2, f3)])]
            else:
                raise ValueError("inception block type %r (the reference knows '3path' and '2path')" % (btype,))
            out.append((base, paths, sum(p[1][-1][3] for p in paths)))
    return out


def upscaler_incep_resnet_layer_specs(filters=64, upscale_factor=4, a=("3path", 5, 3), b=("2path", 10, 7), c=("2path", 5, 3)):
    """make_upscaler_incep_resnet (model.py:445-497) in graph order with the reference's layer names."""
    f = filters
    specs = [("initial/conv/9x9", "conv", dict(kh=9, kw=9, cin=3, cout=f))]
    for base, paths, cat in incep_resnet_blocks(f, a, b, c):
        for _, minis in paths:
            for name, (kh, kw), cin, cout in minis:
                specs += [(name + "/batch_norm", "bn", dict(c=cin)), (name + "/prelu", "prelu", dict(c=cin)),
                          (name + "/%dx%d" % (kh, kw), "conv", dict(kh=kh, kw=kw, cin=cin, cout=cout))]
        specs.append((base + "/final/1x1", "conv", dict(kh=1, kw=1, cin=cat, cout=f)))
    ck = c[2]
    specs += [("prefinal/conv2d", "conv", dict(kh=ck, kw=ck, cin=f, cout=f)), ("prefinal/batch_norm", "bn", dict(c=f))]
    cin = f
    for i in range(int(math.log(upscale_factor, 2))):
        specs.append(("upscaling/%d/block/conv_transp" % i, "convT", dict(kh=ck, kw=ck, cin=cin, cout=256)))
        cin = 256
    specs.append(("final/conv", "conv", dict(kh=9, kw=9, cin=cin, cout=3)))
    return specs


VGG19_CFG = [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512]
VGG19_NAMES = ["block1_conv1", "block1_conv2", "block2_conv1", "block2_conv2", "block3_conv1", "block3_conv2",
               "block3_conv3", "block3_conv4", "block4_conv1", "block4_conv2", "block4_conv3", "block4_conv4",
               "block5_conv1", "block5_conv2", "block5_conv3", "block5_conv4"]


def vgg19_layer_specs():
    specs, cin, i = [], 3, 0
    for c in VGG19_CFG:
        if c == "M":
            continue
        specs.append((VGG19_NAMES[i], "conv", dict(kh=3, kw=3, cin=cin, cout=c)))
        cin, i = c, i + 1
    return specs


DISC_FILTERS = [64, 128, 256, 512, 512, 512, 512, 512, 512]          # make_discriminator_simple_512, model.py:839-873
DISC_THIN_FILTERS = [64, 128, 128, 128, 128, 128, 128, 128, 128]     # make_discriminator_thin_512, model.py:901-939


def discriminator_simple_512_layer_specs(input_shape, filters=None):
    h, w, cin = input_shape
    specs = []
    for i, c in enumerate(filters or DISC_FILTERS):
        specs.append(("discriminator/block_%d/Conv2d" % (i + 1), "conv", dict(kh=3, kw=3, cin=cin, cout=c, stride=1 if i == 0 else 2)))
        specs.append(("discriminator/block_%d/BatchNorm" % (i + 1), "bn", dict(c=c)))
        if i > 0:
            h, w = -(-h // 2), -(-w // 2)
        cin = c
    flat = h * w * cin
    for j, (fi, fo) in enumerate([(flat, 1024), (1024, 32), (32, 1)]):
        specs.append(("discriminator/final/Dense_%d" % (j + 1), "dense", dict(cin=fi, cout=fo)))
        if j < 2:
            specs.append(("discriminator/final/BatchNorm_%d" % (j + 1), "bn", dict(c=fo)))
    return specs


DISC_SPARSE_FILTERS = [64, 128, 256, 256, 256, 256]      # make_discriminator_sparse_512, model.py:967-989
DISC_SPARSE_STRIDES = [1, 3, 3, 3, 3, 3]


def discriminator_sparse_512_layer_specs(input_shape):
    """5x5 padding='valid' blocks (512 -> 508 -> 168 -> 55 -> 17 -> 5 -> 1), Dense 128 / 32 / 1 (model.py:964-1000)."""
    h, w, cin = input_shape
    specs = []
    for i, (c, st) in enumerate(zip(DISC_SPARSE_FILTERS, DISC_SPARSE_STRIDES)):
        specs.append(("discriminator/block_%d/Conv2d" % (i + 1), "conv", dict(kh=5, kw=5, cin=cin, cout=c, stride=st, padding="valid")))
        specs.append(("discriminator/block_%d/BatchNorm" % (i + 1), "bn", dict(c=c)))
        h, w = (h - 5) // st + 1, (w - 5) // st + 1
        if h < 1 or w < 1:
            raise ValueError("make_discriminator_sparse_512 needs inputs of at least 489 x 489, got %s" % (tuple(input_shape),))
        cin = c
    flat = h * w * cin
    for j, (fi, fo) in enumerate([(flat, 128), (128, 32), (32, 1)]):
        specs.append(("discriminator/final/Dense_%d" % (j + 1), "dense", dict(cin=fi, cout=fo)))
        if j < 2:
            specs.append(("discriminator/final/BatchNorm_%d" % (j + 1), "bn", dict(c=fo)))
    return specs


def init_keras_weights(specs, seed=None):
    """Keras initialisers: glorot_uniform kernels, zero biases, BN (1,0,0,1), PReLU alpha 0 (model.py:21; Keras
    default for :276)."""
    rng = np.random.default_rng(seed)
    w = OrderedDict()
    for name, kind, c in specs:
        if kind == "conv":
            w[name] = [glorot_uniform(rng, (c["kh"], c["kw"], c["cin"], c["cout"]), c["kh"] * c["kw"] * c["cin"],
                                      c["kh"] * c["kw"] * c["cout"]), np.zeros(c["cout"], np.float32)]
        elif kind == "convT":
            w[name] = [glorot_uniform(rng, (c["kh"], c["kw"], c["cout"], c["cin"]), c["kh"] * c["kw"] * c["cin"],
                                      c["kh"] * c["kw"] * c["cout"]), np.zeros(c["cout"], np.float32)]
        elif kind == "bn":
            w[name] = [np.ones(c["c"], np.float32), np.zeros(c["c"], np.float32), np.zeros(c["c"], np.float32),
                       np.ones(c["c"], np.float32)]
        elif kind == "prelu":
            w[name] = [np.zeros((1, 1, c["c"]), np.float32)]
        elif kind == "dense":
            w[name] = [glorot_uniform(rng, (c["cin"], c["cout"]), c["cin"], c["cout"]), np.zeros(c["cout"], np.float32)]
    return w


def init_vgg19_weights(seed=4321):
    """He-normal stand-in for the ImageNet weights keras.applications would download (model.py:108); same seeding as
    the CPU oracle so parity runs share weights."""
    rng = np.random.default_rng(seed)
    w, cin, i = OrderedDict(), 3, 0
    for c in VGG19_CFG:
        if c == "M":
            continue
        k = (rng.standard_normal((3, 3, cin, c)) * math.sqrt(2.0 / (9 * cin))).astype(np.float32)
        w[VGG19_NAMES[i]] = [k, (rng.standard_normal(c) * 0.01).astype(np.float32)]
        cin, i = c, i + 1
    return w


def _shapes(kind, c):
    if kind == "conv":
        return [(c["kh"], c["kw"], c["cin"], c["cout"]), (c["cout"],)]
    if kind == "convT":
        return [(c["kh"], c["kw"], c["cout"], c["cin"]), (c["cout"],)]
    if kind == "bn":
        return [(c["c"],)] * 4
    if kind == "prelu":
        return [(1, 1, c["c"])]
    if kind == "dense":
        return [(c["cin"], c["cout"]), (c["cout"],)]
    raise ValueError(kind)


class Net:
    """Flat parameter store + packed 16-bit weight copies shared by the three networks."""

    train_dtype = torch.float16    # see TRAIN_DTYPE; every kernel takes the element type per call

    def __init__(self, specs, weights, seed, device):
        self.on_gpu = torch.cuda.is_available()
        # Without a CUDA device the object still carries structure and weights (shapes, parameter counts, .h5 I/O);
        # anything that computes raises, there is no CPU execution path.
        self.device = torch.device(device) if self.on_gpu else torch.device("cpu")
        self.specs = specs
        host = weights if weights is not None else init_keras_weights(specs, seed)
        # ---- flat layout: trainable tensors first, BatchNorm moving statistics after them
        entries, off = [], 0
        for trainable_pass in (True, False):
            for name, kind, cfg in specs:
                for j, shp in enumerate(_shapes(kind, cfg)):
                    is_moving = kind == "bn" and j >= 2
                    if is_moving == trainable_pass:
                        continue
                    n = int(np.prod(shp))
                    entries.append((name, j, shp, off, n))
                    off += -(-n // ALIGN) * ALIGN
            if trainable_pass:
                self.n_trainable = off
        self.n_flat = off
        flat = np.zeros(self.n_flat, np.float32)
        self.offsets = {}
        for name, j, shp, o, n in entries:
            if name not in host:
                raise KeyError("missing weights for layer %s" % name)
            a = np.asarray(host[name][j], np.float32)
            if tuple(a.shape) != tuple(shp):
                raise ValueError("layer %s weight %d has shape %s, expected %s" % (name, j, a.shape, shp))
            flat[o:o + n] = a.reshape(-1)
            self.offsets[(name, j)] = (o, n, shp)
        self.flat = torch.from_numpy(flat).to(self.device)
        self.weights = OrderedDict()
        for name, kind, cfg in specs:
            self.weights[name] = [self._view(self.flat, name, j) for j in range(len(_shapes(kind, cfg)))]
        self.grad_flat = None
        self.grads = None
        self._pack_sets = {}      # set name -> dict(dtype, idx tensor, views {key: (offset, shape)})
        self._packed = {}         # set name -> {key: tensor view}
        self._pack_src = {}       # set name -> [(key, layer, idx ndarray)]
        self._buffers = {}
        self.layer_events = None  # dict name -> [(start, end) CUDA events] when bench.py enables per-layer timing

    def _view(self, flat, name, j):
        o, n, shp = self.offsets[(name, j)]
        return flat[o:o + n].view(shp)

    # ---------------------------------------------------------------- gradients
    def ensure_grads(self):
        if self.grad_flat is None:
            self.grad_flat = torch.zeros(self.n_trainable, dtype=torch.float32, device=self.device)
            self.grads = OrderedDict()
            for name, kind, cfg in self.specs:
                nt = 2 if kind == "bn" else len(_shapes(kind, cfg))
                self.grads[name] = [self._view(self.grad_flat, name, j) for j in range(nt)]
        return self.grad_flat

    # ---------------------------------------------------------------- packed weights
    def add_pack(self, set_name, key, layer, idx):
        """Registers packed[key] = cast(weights[layer][0].flat[idx]) (idx < 0 -> 0) in pack set `set_name`."""
        self._pack_src.setdefault(set_name, []).append((key, layer, np.asarray(idx)))

    def build_packs(self, dtypes):
        for set_name, items in self._pack_src.items():
            views, chunks, off = {}, [], 0
            for key, layer, idx in items:
                base = self.offsets[(layer, 0)][0]
                g = np.where(idx >= 0, idx.astype(np.int64) + base, -1).astype(np.int32).reshape(-1)
                views[key] = (off, idx.shape)
                chunks.append(g)
                pad = (-g.size) % 64     # keep every packed matrix 128-byte aligned (TMA global address)
                if pad:
                    chunks.append(-np.ones(pad, np.int32))
                off += g.size + pad
            self._pack_sets[set_name] = dict(dtype=dtypes[set_name], views=views, n=off,
                                             idx=torch.from_numpy(np.concatenate(chunks)).to(self.device), buf=None)

    def refresh_packs(self, set_name):
        """Re-derives one set of packed 16-bit weights from the fp32 masters (one launch)."""
        ps = self._pack_sets[set_name]
        ps["buf"] = ops.gather_cast(self.flat, ps["idx"], ps["buf"], ps["dtype"])
        self._packed[set_name] = {k: ps["buf"][o:o + int(np.prod(shp))].view(shp) for k, (o, shp) in ps["views"].items()}
        return self._packed[set_name]

    # ---------------------------------------------------------------- misc
    def host_weights(self):
        return OrderedDict((k, [t.detach().cpu().numpy() for t in ts]) for k, ts in self.weights.items())

    def set_weights(self, host):
        for name, ts in self.weights.items():
            for t, a in zip(ts, host[name]):
                t.copy_(torch.as_tensor(np.asarray(a, np.float32)).to(self.device).view(t.shape))
        self._train_dirty = self._infer_dirty = True
        self._sat_clean = 0          # replaced weights: the next inference pass is range-checked (forward_infer)
        if self.on_gpu:
            self.refresh()

    def count_params(self):
        return int(sum(t.numel() for ts in self.weights.values() for t in ts))

    def trainable_count(self):
        return int(sum(t.numel() for (name, kind, cfg) in self.specs
                       for j, t in enumerate(self.weights[name]) if not (kind == "bn" and j >= 2)))

    def _buf(self, key, shape, dtype):
        b = self._buffers.get(key)
        if b is None or tuple(b.shape) != tuple(shape) or b.dtype != dtype:
            b = torch.empty(shape, dtype=dtype, device=self.device)
            self._buffers[key] = b
        return b

    def _vec(self, key, n, zero=False):
        b = self._buffers.get(key)
        if b is None or b.numel() != n:
            b = torch.zeros(n, dtype=torch.float32, device=self.device)
            self._buffers[key] = b
        elif zero:
            b.zero_()
        return b

    def _timed(self, name, fn):
        if self.layer_events is None:
            return fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = fn()
        e1.record()
        self.layer_events.setdefault(name, []).append((e0, e1))
        return r

    # ---------------------------------------------------------------- fp16 training-storage range guard
    def range_flag(self):
        """One-float device flag the BatchNorm finalize launches of this network raise (None for bf16 storage)."""
        if self.train_dtype != torch.float16:
            return None
        return self.guard_words()[0:1].view(torch.float32)

    def guard_words(self):
        """Two 32-bit device words that stay zero while fp16 storage is safe: [0] the BatchNorm range flag (float bits, raised
        by the finalize launches), [1] the count of saturated values among the activation gradients the backward pass samples.
        The guarded Adam update reads them on the device and drops an update computed from clipped values
        (ops.adam guards=); the host reads them with the step's loss (check_range).  None for bf16 storage."""
        if self.train_dtype != torch.float16:
            return None
        g = self._buffers.get("guard_words")
        if g is None:
            g = self._buffers["guard_words"] = torch.zeros(2, dtype=torch.int32, device=self.device)
        return g

    def probe_grad(self, *tensors):
        """Backward passes call this on a few (small) fp16 activation-gradient tensors: saturated values are counted into
        guard word 1.  No-op for bf16 storage, whose exponent range is fp32's."""
        if self.train_dtype != torch.float16 or not self.on_gpu:
            return
        cnt = self.guard_words()[1:2]
        for t in tensors:
            if t is not None and t.dtype == torch.float16 and t.numel() <= GRAD_PROBE_MAX_ELEMS:
                ops.count_saturated(t, cnt)

    def set_train_dtype(self, dtype):
        """Switches the 16-bit storage type of the training passes (activations, packed weights, gradients)."""
        if dtype not in (torch.float16, torch.bfloat16):
            raise ValueError("training storage is float16 or bfloat16")
        self.train_dtype = dtype
        ps = self._pack_sets.get("train")
        if ps is not None:
            ps["dtype"], ps["buf"] = dtype, None
            if self.on_gpu:
                self.refresh(("train",))

    def check_range(self, extra=(), agree=None):
        """Host side of the guard, once per training step with the step's loss read: reads the two guard words (one small D2H;
        with `extra`, also a saturation count of those stored forward tensors); on a hit warns and drops to bf16 storage, or
        raises (ON_RANGE).  Returns True when the storage type was switched; `self.guard_hit` says whether the guard words
        themselves were raised, i.e. whether the guarded Adam update of this step was dropped on the device.
        `agree` (data-parallel training: local hit -> hit on any replica, a collective) makes all replicas switch together:
        the sampled `extra` count is per replica, and replicas on different storage types would issue different collectives.
        (Guard words the step has already max-reduced over the replicas before its update read them -- `_guard_reduced`,
        set by training.Compiled._adam -- need no second agreement.)"""
        self.guard_hit = False
        guard = self.guard_words()
        if guard is None:
            return False
        cnt = None
        for t in extra:
            if t is not None and t.dtype == torch.float16:
                if cnt is None:
                    cnt = self._buffers.get("range_count")
                    if cnt is None:
                        cnt = self._buffers["range_count"] = torch.zeros(1, dtype=torch.int32, device=self.device)
                ops.count_saturated(t, cnt)
        words = guard.cpu()
        fwd, bwd = int(words[0]) != 0, int(words[1]) != 0
        self.guard_hit = fwd or bwd
        hit = self.guard_hit or (cnt is not None and int(cnt.item()) != 0)
        reduced, self._guard_reduced = getattr(self, "_guard_reduced", False), False
        if agree is not None and (cnt is not None or not reduced):
            hit = agree(hit)
        if not hit:
            return False
        guard.zero_()
        if cnt is not None:
            cnt.zero_()
        what = []
        if fwd:
            what.append("a BatchNorm input channel with RMS >= %g" % RANGE_LIMIT)
        if bwd:
            what.append("saturated activation gradients")
        if not what or (cnt is not None and not self.guard_hit):
            what.append("saturated stored values")
        msg = ("%s: the training passes approached the edge of float16 storage (%s)%s"
               % (type(self).__name__, ", ".join(what), "; this step's weight update was dropped" if self.guard_hit else ""))
        if ON_RANGE == "raise":
            raise FloatingPointError(msg)
        import warnings
        warnings.warn(msg + "; switching this network's training storage to bfloat16", RuntimeWarning)
        self.set_train_dtype(torch.bfloat16)
        return True

    def _need_gpu(self):
        if not self.on_gpu:
            raise RuntimeError("this network runs on hand-written sm_100a kernels only; no CUDA device is available")

    # ---------------------------------------------------------------- train-mode BatchNormalization helpers
    def _bn_forward(self, tape, key, bn_name, z, out, act, alpha=None, leaky=0.0, residual=None, sync=None, stats=None,
                    update_moving=True):
        """Batch statistics of z (+ the other replicas' when `sync` is given), moving-statistics update,
        out = act(BN(z)) + residual.  `stats` = (part, nparts) when the kernel that produced z already wrote the partial
        sums (conv epilogue).  The statistics vectors are allocated per call and owned by the tape: a second forward
        through the same layer before the first one's backward (real / fake batches, train_gan3.py:353) keeps its own."""
        C = z.shape[-1]
        rows = z.numel() // C
        part, nparts = stats if stats is not None else ops.bn_stats_partial(z)
        aff = torch.empty(4 * C, dtype=torch.float32, device=z.device)       # scale | shift | mean | invstd
        gamma, beta, mm, mv = self.weights[bn_name]
        if not update_moving:
            mm = mv = None
        world = sync.world if sync is not None else 1
        peer = sync.peer_for(z.device) if sync is not None else None
        watch = self.range_flag() if z.dtype == torch.float16 else None
        if sync is None or peer is not None:
            ops.bn_fwd_finalize(part, nparts, C, rows * world, gamma, beta, aff, mm, mv, peer=peer, eps=BN_EPS, momentum=BN_MOMENTUM,
                                range_watch=watch, range_limit=RANGE_LIMIT)
        else:                   # replicas that cannot map each other's memory: sums over NCCL, two stage-2 launches
            sums = torch.empty(2 * C, dtype=torch.float32, device=z.device)
            ops.bn_fwd_finalize(part, nparts, C, rows, None, None, None, sums_out=sums)
            sync.all_reduce(sums)
            ops.bn_fwd_finalize(sums, 1, C, rows * world, gamma, beta, aff, mm, mv, eps=BN_EPS, momentum=BN_MOMENTUM,
                                range_watch=watch, range_limit=RANGE_LIMIT)
        ops.affine_act(z, out, aff[:C], aff[C:2 * C], alpha, act, leaky, residual)
        tape[key] = (z, aff, rows * world)
        return out

    def _bn_backward(self, tape, key, bn_name, g, dz, act, alpha=None, leaky=0.0, dalpha=None, sync=None, want_param_grads=True,
                     stats=None):
        """g = gradient w.r.t. act(BN(z)); ACCUMULATES dgamma / dbeta (and dalpha) into the gradient views, returns dz.
        The two reductions go through a per-call scratch (never through the gradient accumulators), so several backward
        passes may add into the same gradient buffer."""
        z, aff, rows = tape[key]
        C = z.shape[-1]
        part, nparts = stats if stats is not None else ops.bn_bwd_partial(z, g, aff, alpha, act, leaky)
        sums = torch.empty(2 * C, dtype=torch.float32, device=z.device)
        dgamma, dbeta = (self.grads[bn_name][0], self.grads[bn_name][1]) if want_param_grads else (None, None)
        da = dalpha if want_param_grads else None
        world = sync.world if sync is not None else 1
        peer = sync.peer_for(z.device) if sync is not None else None
        # dgamma / dbeta come from sums over ALL replicas; the final gradient all-reduce sums the replicas again: 1/world
        if sync is None or peer is not None:
            ops.bn_bwd_finalize(part, nparts, C, sums, dgamma, dbeta, da, 1.0 / world, peer=peer)
        else:
            raw = torch.empty(3 * C, dtype=torch.float32, device=z.device)
            ops.bn_bwd_finalize(part, nparts, C, None, raw_out=raw)
            sync.all_reduce(raw[:2 * C])
            ops.bn_bwd_finalize(raw, 1, C, sums, dgamma, dbeta, da, 1.0 / world)
        ops.bn_bwd_apply(z, g, aff[:C], aff[C:2 * C], alpha, act, leaky, aff[2 * C:3 * C], aff[3 * C:], sums[:C], sums[C:], rows, dz)
        return dz


class UpscalerOrigNet(Net):
    """SRGAN generator: 9x9 conv + PReLU -> N x (conv-BN-PReLU-conv-BN-add) -> conv-BN-add(skip) ->
    log2(factor) x (ConvT s2 + LeakyReLU 0.2) -> 9x9 conv -> tanh."""

    input_name = "initial/input"
    writes_out = True      # forward_infer(x, out=...) writes its result into a caller-provided tensor (engine.predict)
    # Saturation guard of the fp16 inference path: the conv epilogues store with a saturating conversion (+-65504), so the
    # FIRST forward pass after every weight change (and any pass with check_saturation=True) counts, layer by layer, the
    # stored activations that sit at that edge.  A non-zero count is never silent: "fallback" warns, switches this
    # network's inference storage to bf16 (fp32's range, 3 fewer significand bits) and reruns the batch; "raise" raises
    # FloatingPointError.  `saturation_report` keeps the last per-layer counts.
    on_saturation = "fallback"
    SAT_CLEAN_STREAK, SAT_CHECK_EVERY = 4, 16
    # Frames per trunk chunk of the inference pass: None = the whole batch in one launch per layer (see _forward_infer).
    TRUNK_CHUNK = None
    extra_layers = [("prefinal/tanh", "add"), ("final/tanh", "activation")]

    # graph variant (overridden by UpscalerSkipConNet)
    up_filters, up_kernel, prefinal_kernel, concat_input, supports_training = 256, None, None, False, True

    def __init__(self, kernel_size=5, filters=64, upscale_factor=4, res_block_num=16, weights=None, device="cuda",
                 seed=None, infer_dtype=torch.float16, train_dtype=None):
        if filters != 64:
            raise ValueError("the B200 path is built for filters=64 (reference default, model.py:267)")
        self.kernel_size, self.filters = kernel_size, filters
        self.upscale_factor, self.res_block_num = upscale_factor, res_block_num
        self.upscale_times = int(math.log(upscale_factor, 2))
        # 16-bit storage type of the inference path (activations + packed weights): fp16 keeps 11 significand bits,
        # which is what holds max-abs error under 2e-2 across 37 layers; bf16 is accepted and ~8x noisier.
        self.infer_dtype = infer_dtype
        self.train_dtype = train_dtype or TRAIN_DTYPE
        self.up_kernel = self.up_kernel or kernel_size
        self.prefinal_kernel = self.prefinal_kernel or kernel_size
        super().__init__(self._layer_specs(), weights, seed, device)
        self._infer = {}
        self._final_rows = False
        self._sat_pending, self.saturation_report = True, {}
        self._build_index_maps()
        if self.on_gpu:
            self.refresh()

    def _layer_specs(self):
        return upscaler_orig_layer_specs(self.kernel_size, self.filters, self.upscale_factor, self.res_block_num, self.up_filters,
                                         self.up_kernel, self.prefinal_kernel, self.concat_input)

    # ---------------------------------------------------------------- weights
    def _build_index_maps(self):
        k = self.kernel_size
        pad64 = lambda n: -(-n // 64) * 64
        self._final_chan_map, self._convT_wgrad_idx, self._convT_bias_idx = None, {}, {}
        for name, kind, c in self.specs:
            if kind == "conv":
                if name == "final/conv" and self.concat_input:
                    # our tensor: channels [0, F) = upsampled features, [F, F+3) = resized input, rest zero padding;
                    # Keras: Concatenate([resized input (3), features (F)]) (model.py:358)
                    F = self.up_filters
                    cm = -np.ones(pad64(F + 3), np.int64)
                    cm[:F] = 3 + np.arange(F)
                    cm[F:F + 3] = np.arange(3)
                    idx = ops.conv_scatter_pack_index(c["kh"], c["kw"], c["cin"], c["cout"], chan_map=cm)
                    self._final_rows = True
                    self._final_chan_map = cm
                elif name == "final/conv" and c["kh"] == 9 and c["kw"] == 9 and c["cout"] == 3 and c["cin"] % 64 == 0 and c["cin"] <= 256:
                    idx = ops.conv_scatter_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
                    self._final_rows = True
                elif c["cin"] < 64:
                    idx = ops.conv_rowfold_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
                else:
                    idx = ops.conv_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
            elif kind == "convT":
                idx = ops.convT_pack_index(c["kh"], c["kw"], c["cin"], c["cout"], pad64(c["cin"]), pad64(c["cout"]))
            else:
                continue
            self.add_pack("infer", name, name, idx)
            if not self.supports_training:
                continue
            self.add_pack("train", name, name, idx)
            # data-gradient layouts (training only)
            if kind == "conv" and c["cin"] >= 64 and name != "final/conv":
                self.add_pack("train", name + "#dgrad", name, ops.conv_dgrad_pack_index(c["kh"], c["kw"], c["cin"], c["cout"]))
            elif kind == "convT":
                self.add_pack("train", name + "#dgrad", name, ops.convT_dgrad_pack_index(c["kh"], c["kw"], c["cin"], c["cout"],
                                                                                         pad64(c["cin"]), pad64(c["cout"])))
                if c["cout"] % 64 or c["cin"] % 64:       # weight / bias gradients come out in the padded layout
                    self._convT_bias_idx[name] = torch.arange(c["cout"], dtype=torch.int32, device=self.device)
                    self._convT_wgrad_idx[name] = torch.from_numpy(ops.padded_gather_index(
                        (c["kh"], c["kw"], c["cout"], c["cin"]), (c["kh"], c["kw"], pad64(c["cout"]), pad64(c["cin"])))).to(self.device)
            elif name == "final/conv":
                self.add_pack("train", name + "#dgrad", name, ops.final_dgrad_pack_index(c["kh"], c["kw"], c["cin"], c["cout"],
                                                                                         chan_map=self._final_chan_map))
        self.build_packs({"infer": self.infer_dtype, "train": self.train_dtype} if self.supports_training else
                         {"infer": self.infer_dtype})
        c0 = self.specs[0][2]
        self._head_wgrad_idx = torch.from_numpy(ops.rowfold_wgrad_gather_index(c0["kh"], c0["kw"], c0["cin"], c0["cout"])).to(self.device)
        cf = self.specs[-1][2]
        if self.supports_training:
            self._final_wgrad_idx = torch.from_numpy(ops.final_wgrad_gather_index(cf["kh"], cf["kw"], cf["cin"], cf["cout"],
                                                                                  chan_map=self._final_chan_map)).to(self.device)

    def refresh(self, sets=("infer",)):
        """Re-derive packed 16-bit weights (and, for inference, the folded BatchNorm epilogues) from the fp32 masters."""
        for s in sets:
            self.refresh_packs(s)
            setattr(self, "_%s_dirty" % s, False)
        if "infer" not in sets:
            return
        self._sat_pending = True
        w = self.weights
        nfold = 2 * self.res_block_num + 1
        fold = self._vec("fold", 2 * 64 * nfold)
        inf, slot = {}, [0]

        def fold_bn(conv, bn):
            i = slot[0]
            slot[0] += 1
            scale, shift = fold[2 * 64 * i:2 * 64 * i + 64], fold[2 * 64 * i + 64:2 * 64 * (i + 1)]
            gamma, beta, mean, var = w[bn]
            ops.bn_fold(gamma, beta, mean, var, w[conv][1], scale, shift, BN_EPS)
            return scale, shift

        inf["initial/conv"] = (None, w["initial/conv"][1], w["initial/prelu"][0].reshape(-1))
        for i in range(self.res_block_num):
            n = "res_block/%d" % i
            s, b = fold_bn(n + "/conv_pre", n + "/batch_norm_pre")
            inf[n + "/conv_pre"] = (s, b, w[n + "/prelu"][0].reshape(-1))
            s, b = fold_bn(n + "/conv_post", n + "/batch_norm_post")
            inf[n + "/conv_post"] = (s, b, None)
        s, b = fold_bn("prefinal/conv2d", "prefinal/batch_norm")
        inf["prefinal/conv2d"] = (s, b, None)
        for i in range(self.upscale_times):
            n = "upscaling/%d/block/conv_transp" % i
            bias = w[n][1]
            if bias.numel() % 64:                       # padded output channels: zero bias -> LeakyReLU(0) = 0
                pb = self._vec("bias_pad/" + n, -(-bias.numel() // 64) * 64, zero=True)
                pb[:bias.numel()].copy_(bias)
                bias = pb
            inf[n] = (None, bias, None)
        inf["final/conv"] = (None, w["final/conv"][1], None)
        self._infer = inf

    def output_shape_for(self, input_shape):
        h, w, c = input_shape
        f = 2 ** self.upscale_times
        return (None if h is None else h * f, None if w is None else w * f, 3)

    # ---------------------------------------------------------------- inference
    def forward_infer(self, x, out=None, check_saturation=None):
        """x: (B,H,W,3) float32 CUDA in [-1,1] -> (B, H*f, W*f, 3) float32.  BatchNorm uses moving statistics folded
        into the conv epilogues (Keras learning_phase=0, i.e. Model.predict).  `check_saturation`: None = only on the
        first pass after a weight change; see `on_saturation`."""
        self._need_gpu()
        if getattr(self, "_infer_dirty", False):     # fp32 masters changed (training step / set_weights)
            self.refresh(("infer",))
        B, H, W, C = x.shape
        if C != 3 or x.dtype != torch.float32:
            raise ValueError("expected (B,H,W,3) float32 input, got %s %s" % (tuple(x.shape), x.dtype))
        check = self._sat_pending if check_saturation is None else bool(check_saturation)
        if check and check_saturation is None:
            # weights that change every iteration (predict inside a training loop, train_gan.py:305): once SAT_CLEAN_STREAK
            # consecutive weight versions passed the check, only every SAT_CHECK_EVERY-th version is checked (36 extra launches
            # and a device -> host read per checked pass)
            self._sat_version = getattr(self, "_sat_version", 0) + 1
            if getattr(self, "_sat_clean", 0) >= self.SAT_CLEAN_STREAK and self._sat_version % self.SAT_CHECK_EVERY:
                check, self._sat_pending = False, False
        counters = self._sat_counters(check)
        y = self._forward_infer(x, out, counters)
        if check:
            counts = counters.cpu().numpy()          # the one host read of a checked pass
            self._sat_pending = False
            self.saturation_report = {n: int(c) for n, c in zip(self._sat_names(), counts) if c}
            self._sat_clean = 0 if self.saturation_report else getattr(self, "_sat_clean", 0) + 1
            if self.saturation_report:
                msg = ("%s: %d stored activations reached the edge of %s in %s (first layer: %s)" % (
                    type(self).__name__, int(counts.sum()), str(self.infer_dtype).replace("torch.", ""),
                    "%d layers" % len(self.saturation_report), next(iter(self.saturation_report))))
                if self.on_saturation == "raise" or self.infer_dtype != torch.float16:
                    raise FloatingPointError(msg)
                import warnings
                warnings.warn(msg + "; switching this network's inference storage to bfloat16 and rerunning", RuntimeWarning)
                self.set_infer_dtype(torch.bfloat16)
                return self.forward_infer(x, out, check_saturation=True)
        return y

    def set_infer_dtype(self, dtype):
        """Switches the 16-bit storage type of the inference path (activations + packed weights)."""
        if dtype not in (torch.float16, torch.bfloat16):
            raise ValueError("inference storage is float16 or bfloat16")
        self.infer_dtype = dtype
        ps = self._pack_sets["infer"]
        ps["dtype"], ps["buf"] = dtype, None
        if self.on_gpu:
            self.refresh(("infer",))

    def trunk_chunk_for(self, B, H, W):
        c = self.TRUNK_CHUNK
        return max(1, B) if not c or c >= B else max(1, int(c))

    def _sat_names(self):
        names = ["initial/conv"]
        for i in range(self.res_block_num):
            names += ["res_block/%d/conv_pre" % i, "res_block/%d/conv_post" % i]
        names.append("prefinal/conv2d")
        names += ["upscaling/%d/block/conv_transp" % i for i in range(self.upscale_times)]
        return names

    def _sat_counters(self, check):
        if not check:
            return None
        c = self._buffers.get("sat_counters")
        n = len(self._sat_names())
        if c is None or c.numel() != n:
            c = self._buffers["sat_counters"] = torch.zeros(n, dtype=torch.int32, device=self.device)
        else:
            c.zero_()
        return c

    def _forward_infer(self, x, out, counters):
        B, H, W, C = x.shape
        k, f, dt = self.kernel_size, self.filters, self.infer_dtype
        inf, pk = self._infer, self._packed["infer"]
        slot = [0]

        def chk(t):          # saturation guard: count this layer's stored activations at the edge of the 16-bit type
            if counters is not None:
                ops.count_saturated(t, counters[slot[0]:slot[0] + 1])
                slot[0] += 1
            return t
        # head: fold the 9 horizontal taps of the 3-channel input into 27 (of 64) channels, then a 9x1 conv
        T = self._timed
        xp = T("initial/fold", lambda: ops.fold_taps(x, [(0, dx - 4) for dx in range(9)],
                                                     out=self._buf("fold", (B, H, W, 64), dt), dtype=dt))
        s, b, a = inf["initial/conv"]
        skip = chk(T("initial/conv", lambda: ops.conv2d(xp, pk["initial/conv"], f, 9, 1,
                                                        out=self._buf("skip", (B, H, W, f), dt), scale=s, shift=b, alpha=a,
                                                        act=ops.ACT_PRELU)))
        # The residual trunk walks the batch in chunks of `trunk_chunk` frames (None = whole batch): a 64-channel trunk tensor
        # of 8 frames is 133 MB -- each layer then streams its input, residual and output through HBM -- while chunks whose three
        # live tensors fit the 126 MB L2 keep the layer chain on-chip at the price of more launches (see TRUNK_CHUNK).
        ping = [self._buf("t%d" % i, (B, H, W, f), dt) for i in range(3)]
        kp, ku = self.prefinal_kernel, self.up_kernel
        upc = -(-self.up_filters // 64) * 64           # 224 -> 256: padded channels carry zeros
        chunk = self.trunk_chunk_for(B, H, W)
        slot0, pi_out = slot[0], 0
        for b0 in range(0, B, chunk):
            b1 = min(B, b0 + chunk)
            slot[0] = slot0                            # the saturation counters of a layer add up over the chunks
            skip_c, ping_c = skip[b0:b1], [t[b0:b1] for t in ping]
            m = skip_c
            for i in range(self.res_block_num):
                n = "res_block/%d" % i
                free = [t for t in ping_c if t is not m]
                s, b, a = inf[n + "/conv_pre"]
                h = chk(T("res_block/*/conv_pre", lambda: ops.conv2d(m, pk[n + "/conv_pre"], f, k, k, out=free[0], scale=s, shift=b,
                                                                     alpha=a, act=ops.ACT_PRELU)))
                s, b, _ = inf[n + "/conv_post"]
                m = chk(T("res_block/*/conv_post", lambda: ops.conv2d(h, pk[n + "/conv_post"], f, k, k, out=free[1], scale=s,
                                                                      shift=b, residual=m)))
            free = [t for t in ping_c if t is not m]
            s, b, _ = inf["prefinal/conv2d"]
            m = chk(T("prefinal/conv2d", lambda: ops.conv2d(m, pk["prefinal/conv2d"], 64, kp, kp, out=free[0], scale=s, shift=b,
                                                            residual=skip_c)))
            pi_out = [j for j, t in enumerate(ping_c) if t is m][0]
        m = ping[pi_out]
        h, w_ = H, W
        for i in range(self.upscale_times):
            n = "upscaling/%d/block/conv_transp" % i
            _, b, _ = inf[n]
            ob = self._buf("up%d" % i, (B, 2 * h, 2 * w_, upc), dt)
            m = chk(T(n, lambda: ops.conv2d_transpose_s2(m, pk[n], upc, ku, ku, out=ob, shift=b, act=ops.ACT_LEAKY, leaky=0.2)))
            h, w_ = 2 * h, 2 * w_
        if self.concat_input:      # Concatenate([bilinear-resized input, features]) = three of the padding channels
            T("final/resize_input", lambda: ops.resize_bilinear_into(x, m, self.up_filters, 2 ** self.upscale_times))
        _, b, _ = inf["final/conv"]
        if out is None:
            out = torch.empty((B, h, w_, 3), dtype=torch.float32, device=self.device)
        if self._final_rows:
            return T("final/conv", lambda: ops.conv2d_scatter(m, pk["final/conv"], 3, 9, 9, bias=b, act=ops.ACT_TANH, out=out))
        return T("final/conv", lambda: ops.conv2d(m, pk["final/conv"], 3, 9, 9, out=out, shift=b, act=ops.ACT_TANH))

    # ---------------------------------------------------------------- training
    def forward_train(self, x, sync=None, update_moving=True):
        """Training-phase forward (BatchNorm on batch statistics, moving statistics updated).  Returns
        (y fp32 (B,4H,4W,3), tape).  Packed 'train' weights must be current (refresh(('train',)))."""
        self._need_gpu()
        if not self._final_rows:
            raise RuntimeError("training needs the row-streaming final conv (Cin <= 256)")
        B, H, W, _ = x.shape
        k, f, dt = self.kernel_size, self.filters, self.train_dtype
        pk, w = self._packed["train"], self.weights
        tape = {}
        new = lambda shape: torch.empty(shape, dtype=dt, device=self.device)
        xp = ops.fold_taps(x, [(0, dx - 4) for dx in range(9)], out=new((B, H, W, 64)), dtype=dt)
        u0 = ops.conv2d(xp, pk["initial/conv"], f, 9, 1, out=new((B, H, W, f)), shift=w["initial/conv"][1])
        a0 = w["initial/prelu"][0].reshape(-1)
        skip = ops.affine_act(u0, new((B, H, W, f)), alpha=a0, act=ops.ACT_PRELU)
        tape["head"] = (xp, u0)
        m = skip
        for i in range(self.res_block_num):
            n = "res_block/%d" % i
            # the BatchNorm statistics of each conv output come out of the conv's own epilogue (stats_part): no extra pass
            st1 = ops.new_stats_part(f, self.device)
            z1 = ops.conv2d(m, pk[n + "/conv_pre"], f, k, k, out=new((B, H, W, f)), shift=w[n + "/conv_pre"][1], stats_part=st1[0])
            h = self._bn_forward(tape, n + "/pre", n + "/batch_norm_pre", z1, new((B, H, W, f)), ops.ACT_PRELU,
                                 alpha=w[n + "/prelu"][0].reshape(-1), sync=sync, update_moving=update_moving, stats=st1)
            st2 = ops.new_stats_part(f, self.device)
            z2 = ops.conv2d(h, pk[n + "/conv_post"], f, k, k, out=new((B, H, W, f)), shift=w[n + "/conv_post"][1], stats_part=st2[0])
            mo = self._bn_forward(tape, n + "/post", n + "/batch_norm_post", z2, new((B, H, W, f)), ops.ACT_NONE, residual=m,
                                  sync=sync, update_moving=update_moving, stats=st2)
            tape[n] = (m, h)
            m = mo
        kp, ku = self.prefinal_kernel, self.up_kernel
        upc = -(-self.up_filters // 64) * 64
        stp = ops.new_stats_part(64, self.device)
        zp = ops.conv2d(m, pk["prefinal/conv2d"], 64, kp, kp, out=new((B, H, W, 64)), shift=w["prefinal/conv2d"][1],
                        stats_part=stp[0])
        t = self._bn_forward(tape, "prefinal", "prefinal/batch_norm", zp, new((B, H, W, 64)), ops.ACT_NONE, residual=skip, sync=sync,
                             update_moving=update_moving, stats=stp)
        tape["prefinal/in"] = m
        ups, h_, w_ = [t], H, W
        for i in range(self.upscale_times):
            n = "upscaling/%d/block/conv_transp" % i
            bias = w[n][1]
            if bias.numel() != upc:                     # padded output channels: zero bias -> LeakyReLU(0) = 0
                pb = self._vec("bias_pad/" + n, upc, zero=True)
                ops.cast_scale(bias, pb[:bias.numel()], 1.0)
                bias = pb
            u = ops.conv2d_transpose_s2(ups[-1], pk[n], upc, ku, ku, out=new((B, 2 * h_, 2 * w_, upc)), shift=bias,
                                        act=ops.ACT_LEAKY, leaky=0.2)
            ups.append(u)
            h_, w_ = 2 * h_, 2 * w_
        if self.concat_input:       # Concatenate([bilinear-resized input, features]): three of the padding channels
            ops.resize_bilinear_into(x, ups[-1], self.up_filters, 2 ** self.upscale_times)
        tape["ups"] = ups
        y = ops.conv2d_scatter(ups[-1], pk["final/conv"], 3, 9, 9, bias=w["final/conv"][1], act=ops.ACT_TANH)
        tape["y"] = y
        return y, tape

    def backward(self, tape, dz_final, sync=None, on_ready=None):
        """dz_final: fp32 (B,4H,4W,3) gradient w.r.t. the final conv's pre-tanh output (ops.tanh_bwd of the loss
        gradients).  Accumulates every weight gradient into the flat gradient buffer.  `on_ready(offset)` is called
        whenever every gradient at flat offsets >= offset is final (the flat layout follows the graph order, backward
        walks it from the tail), so a data-parallel caller can all-reduce that slice while the rest still runs."""
        self.ensure_grads()
        ready = on_ready if on_ready is not None else (lambda off: None)
        k, f, dt = self.kernel_size, self.filters, self.train_dtype
        pk, w, g = self._packed["train"], self.weights, self.grads
        new = lambda shape: torch.empty(shape, dtype=dt, device=self.device)
        ups = tape["ups"]
        B, HH, WW, _ = dz_final.shape
        # ---- final 9x9 conv: fold the 81 shifted copies of the 3-channel gradient into 243 (of 256) channels; then the
        #      data gradient is a 1x1 conv 256->256 and the weight gradient a plain GEMM over pixels
        ops.colsum_f32(dz_final, g["final/conv"][1])
        F = new((B, HH, WW, 256))
        for q in range(4):
            taps = [(4 - t // 9, 4 - t % 9) for t in range(21 * q, min(21 * q + 21, 81))]
            ops.fold_taps(dz_final, taps, out=F, dtype=dt, out_c0=64 * q)
        raw = self._buf("final_raw", (256 * 256,), torch.float32)
        raw.zero_()
        ops.conv2d_wgrad(ups[-1], F, 1, 1, 1, raw)
        ops.gather_f32(raw, self._final_wgrad_idx, g["final/conv"][0], accumulate=True)
        kp, ku = self.prefinal_kernel, self.up_kernel
        upc = ups[-1].shape[-1]
        # (padding / resized-input channels of `d` meet zero rows of the packed ConvT weights below: their values are moot)
        d = ops.conv2d(F, pk["final/conv#dgrad"], upc, 1, 1, out=new((B, HH, WW, upc)), gate=ups[-1], gate_slope=0.2)
        del F
        # ---- upsampling blocks (Conv2DTranspose stride 2 + LeakyReLU 0.2), last to first
        for i in reversed(range(self.upscale_times)):
            n = "upscaling/%d/block/conv_transp" % i
            xin = ups[i]
            cin = xin.shape[-1]
            if n in self._convT_wgrad_idx:              # padded channels: raw gradients in the padded layout, then gather
                braw = self._vec("bias_raw/" + n, upc, zero=True)
                ops.colstats(d, braw)
                ops.gather_f32(braw, self._convT_bias_idx[n], g[n][1], accumulate=True)
                raw = self._buf("convT_raw/" + n, (ku * ku * upc * cin,), torch.float32)
                raw.zero_()
                ops.conv2d_wgrad(d, xin, ku, ku, 2, raw)
                ops.gather_f32(raw, self._convT_wgrad_idx[n], g[n][0], accumulate=True)
            else:
                ops.colstats(d, g[n][1])
                ops.conv2d_wgrad(d, xin, ku, ku, 2, g[n][0])
            if i > 0:
                d = ops.conv2d(d, pk[n + "#dgrad"], cin, ku, ku, stride=2, out=new(tuple(xin.shape)), gate=xin, gate_slope=0.2)
            else:
                d = ops.conv2d(d, pk[n + "#dgrad"], cin, ku, ku, stride=2, out=new(tuple(xin.shape)))
        dt_ = d      # gradient w.r.t. t = skip + BN(prefinal conv): flows into both the trunk and the head skip
        self.probe_grad(dt_)
        ready(self.offsets[("upscaling/0/block/conv_transp", 0)][0])
        # ---- prefinal conv + BN
        dzp = self._bn_backward(tape, "prefinal", "prefinal/batch_norm", dt_, new(tuple(dt_.shape)), ops.ACT_NONE, sync=sync)
        m_in = tape["prefinal/in"]
        ops.conv2d_wgrad(m_in, dzp, kp, kp, 1, g["prefinal/conv2d"][0])
        gm = ops.conv2d(dzp, pk["prefinal/conv2d#dgrad"], f, kp, kp, out=new(tuple(m_in.shape)))
        # ---- residual blocks, last to first; gm is the gradient w.r.t. each block's output (the residual stream)
        for i in reversed(range(self.res_block_num)):
            n = "res_block/%d" % i
            m_in, h = tape[n]
            dz2 = self._bn_backward(tape, n + "/post", n + "/batch_norm_post", gm, new(tuple(gm.shape)), ops.ACT_NONE, sync=sync)
            ops.conv2d_wgrad(h, dz2, k, k, 1, g[n + "/conv_post"][0])
            dh = ops.conv2d(dz2, pk[n + "/conv_post#dgrad"], f, k, k, out=dz2.new_empty(dz2.shape))
            dz1 = self._bn_backward(tape, n + "/pre", n + "/batch_norm_pre", dh, dz2, ops.ACT_PRELU,
                                    alpha=w[n + "/prelu"][0].reshape(-1), dalpha=g[n + "/prelu"][0].reshape(-1), sync=sync)
            ops.conv2d_wgrad(m_in, dz1, k, k, 1, g[n + "/conv_pre"][0])
            gm = ops.conv2d(dz1, pk[n + "/conv_pre#dgrad"], f, k, k, out=dh, residual=gm)
            if i == self.res_block_num // 2:
                ready(self.offsets[(n + "/conv_pre", 0)][0])
        # ---- head: PReLU backward (dbias = sum dy, dalpha), then the weight gradient of the row-folded 9x1 conv
        xp, u0 = tape["head"]
        a0 = w["initial/prelu"][0].reshape(-1)
        gm = ops.affine_act(gm, new(tuple(gm.shape)), residual=dt_)     # the skip connection joins here: trunk + upsampling path
        self.probe_grad(gm)
        ops.bn_bwd_reduce(u0, gm, None, None, a0, ops.ACT_PRELU, 0.0, None, None, g["initial/conv"][1], None,
                          g["initial/prelu"][0].reshape(-1))
        dz0 = ops.bn_bwd_apply(u0, gm, None, None, a0, ops.ACT_PRELU, 0.0, None, None, None, None, 0, new(tuple(u0.shape)))
        raw = self._buf("head_raw", (9 * 64 * 64,), torch.float32)
        raw.zero_()
        ops.conv2d_wgrad(xp, dz0, 9, 1, 1, raw)
        ops.gather_f32(raw, self._head_wgrad_idx, g["initial/conv"][0], accumulate=True)
        ready(0)


class UpscalerSkipConNet(UpscalerOrigNet):
    """make_upscaler_skip_con (model.py:332-363): the `orig` trunk, a 3x3 after-res conv, two 3x3 Conv2DTranspose blocks
    with 224 filters, and a final 9x9 conv over Concatenate([bilinear x4 input, features]) (227 channels).  The reference
    builds it with unnamed (hence duplicate-named) blocks, which Keras rejects; the layers carry make_upscaler_orig's names
    here.  On the B200 path the 224 channels are padded to 256 with zero weights (LeakyReLU(0) = 0), the resized input is
    written into three of the padding channels, and the final conv's packed weights are permuted accordingly -- so every
    kernel of the `orig` path is reused, in training too (weight gradients of the padded layers are produced in the padded
    layout and gathered back into the Keras one)."""

    up_filters, up_kernel, prefinal_kernel, concat_input, supports_training = 224, 3, 3, True, True


class UpscalerAttentionNet(UpscalerOrigNet):
    """make_upscaler_attention (model.py:296-328): the `orig` trunk with every residual block's input gated by
    sigmoid(conv(network input)) (residual_block_attention, :30-48) and upsampling blocks that gate their input by
    sigmoid(conv(Concatenate([nearest, bilinear] resize of the network input))), run Conv2DTranspose(128, k, stride 2) +
    LeakyReLU(0.2) and add Conv2DTranspose(128, scale + 1, stride scale)(atanh(0.99999 * network input))
    (upsampling_block_attention, :78-98).  Inference path (Model.predict / load_model / save); the training step of this
    variant is not built.
    On the B200 path: the 16 residual-block attention convs all read the same 3-channel image, so they run as ONE
    convolution with 16 x 64 = 1024 output channels over the row-folded input (kw taps x 3 colours in 64 channels, the head
    conv's trick) -- a full-width tensor-core GEMM instead of sixteen N = 64 ones; a block takes its 64-channel slice in the
    gate kernel (b2_gate_mul).  The atanh / stride-`scale` Conv2DTranspose of the 3-channel input is write-bound
    (b2_atanh_conv2d_transpose_small) and enters the block's Conv2DTranspose as the epilogue's residual operand."""

    supports_training = False
    extra_layers = [("after_res/add", "add"), ("final/tanh", "activation")]

    def _layer_specs(self):
        return upscaler_attention_layer_specs(self.kernel_size, self.filters, self.upscale_factor, self.res_block_num)

    def _build_index_maps(self):
        k, f = self.kernel_size, self.filters
        if k * 6 > 64:
            raise ValueError("make_upscaler_attention on the B200 path: kernel_size <= 10 (row fold of the 6-channel attention input)")
        self._final_chan_map, self._convT_wgrad_idx, self._convT_bias_idx = None, {}, {}
        self._final_rows = True
        off0 = None
        att_idx = []
        for name, kind, c in self.specs:
            if name.endswith("to_add_input_conv_transp"):
                continue                                   # read as fp32 masters by b2_atanh_conv2d_transpose_small
            if kind == "conv":
                if name == "final/conv":
                    idx = ops.conv_scatter_pack_index(9, 9, c["cin"], 3)
                elif name.startswith("res_block/") and name.endswith("/attention"):
                    o = self.offsets[(name, 0)][0]
                    off0 = o if off0 is None else off0
                    ix = ops.conv_rowfold_pack_index(k, k, 3, f).astype(np.int64)
                    att_idx.append(np.where(ix >= 0, ix + (o - off0), -1))
                    continue
                elif c["cin"] < 64:
                    idx = ops.conv_rowfold_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
                else:
                    idx = ops.conv_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
            elif kind == "convT":
                idx = ops.convT_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
            else:
                continue
            self.add_pack("infer", name, name, idx)
        if att_idx:      # [res_block_num * 64][k][64]: all residual-block attention convs as one wide convolution
            self.add_pack("infer", "res_block/*/attention", "res_block/0/attention", np.concatenate(att_idx, axis=0).astype(np.int32))
        self.build_packs({"infer": self.infer_dtype})

    def refresh(self, sets=("infer",)):
        for s in sets:
            self.refresh_packs(s)
            setattr(self, "_%s_dirty" % s, False)
        if "infer" not in sets:
            return
        self._sat_pending = True
        w, f, R = self.weights, self.filters, self.res_block_num
        fold = self._vec("fold", 2 * 64 * (2 * R + 1))
        inf, slot = {}, [0]

        def fold_bn(conv, bn):
            i = slot[0]
            slot[0] += 1
            scale, shift = fold[2 * 64 * i:2 * 64 * i + 64], fold[2 * 64 * i + 64:2 * 64 * (i + 1)]
            gamma, beta, mean, var = w[bn]
            ops.bn_fold(gamma, beta, mean, var, w[conv][1], scale, shift, BN_EPS)
            return scale, shift

        inf["initial/conv"] = (None, w["initial/conv"][1], w["initial/prelu"][0].reshape(-1))
        att_bias = self._vec("att_bias", f * R)
        for i in range(R):
            n = "res_block/%d" % i
            att_bias[f * i:f * (i + 1)].copy_(w[n + "/attention"][1])
            s, b = fold_bn(n + "/conv_pre", n + "/batch_norm_pre")
            inf[n + "/conv_pre"] = (s, b, w[n + "/prelu"][0].reshape(-1))
            s, b = fold_bn(n + "/conv_post", n + "/batch_norm_post")
            inf[n + "/conv_post"] = (s, b, None)
        inf["res_block/*/attention"] = (None, att_bias, None)
        s, b = fold_bn("after_res/conv", "after_res/batch_norm")
        inf["after_res/conv"] = (s, b, None)
        self._infer = inf

    def _sat_names(self):
        names = ["initial/conv", "res_block/*/attention"]
        for i in range(self.res_block_num):
            names += ["res_block/%d/conv_pre" % i, "res_block/%d/conv_post" % i]
        names.append("after_res/conv")
        for i in range(self.upscale_times):
            names += ["upscaling/%d/block/attention" % i, "upscaling/%d/block/to_add_input_conv_transp" % i,
                      "upscaling/%d/block/conv_transp" % i]
        return names

    def forward_train(self, *a, **k):
        raise NotImplementedError("make_upscaler_attention: the inference path is built on the B200 kernels; its training step "
                                  "is not (use make_upscaler_orig or make_upscaler_skip_con for training)")

    backward = forward_train

    def _forward_infer(self, x, out, counters):
        B, H, W, C = x.shape
        k, f, dt, R = self.kernel_size, self.filters, self.infer_dtype, self.res_block_num
        inf, pk, w = self._infer, self._packed["infer"], self.weights
        slot = [0]

        def chk(t):
            if counters is not None:
                ops.count_saturated(t, counters[slot[0]:slot[0] + 1])
                slot[0] += 1
            return t
        T = self._timed
        row_taps = lambda kk: [(0, dx - (kk - 1) // 2) for dx in range(kk)]
        xp = T("initial/fold", lambda: ops.fold_taps(x, row_taps(9), out=self._buf("fold", (B, H, W, 64), dt), dtype=dt))
        s, b, a = inf["initial/conv"]
        skip = chk(T("initial/conv", lambda: ops.conv2d(xp, pk["initial/conv"], f, 9, 1, out=self._buf("skip", (B, H, W, f), dt),
                                                        scale=s, shift=b, alpha=a, act=ops.ACT_PRELU)))
        # every residual block's attention map in one convolution: (B,H,W, R*64) pre-sigmoid
        xa = T("res_block/*/attention", lambda: ops.fold_taps(x, row_taps(k), out=self._buf("fold_att", (B, H, W, 64), dt), dtype=dt))
        att = chk(T("res_block/*/attention", lambda: ops.conv2d(xa, pk["res_block/*/attention"], f * R, k, 1,
                                                                out=self._buf("att", (B, H, W, f * R), dt),
                                                                shift=inf["res_block/*/attention"][1])))
        m = skip
        ping = [self._buf("t%d" % i, (B, H, W, f), dt) for i in range(3)]
        gated = self._buf("gated", (B, H, W, f), dt)
        for i in range(R):
            n = "res_block/%d" % i
            free = [t for t in ping if t is not m]
            T("res_block/*/attention_multiply", lambda: ops.gate_mul(att, m, gated, att_c0=f * i))
            s, b, a = inf[n + "/conv_pre"]
            h = chk(T("res_block/*/conv_pre", lambda: ops.conv2d(gated, pk[n + "/conv_pre"], f, k, k, out=free[0], scale=s, shift=b,
                                                                 alpha=a, act=ops.ACT_PRELU)))
            s, b, _ = inf[n + "/conv_post"]
            m = chk(T("res_block/*/conv_post", lambda: ops.conv2d(h, pk[n + "/conv_post"], f, k, k, out=free[1], scale=s,
                                                                  shift=b, residual=m)))
        free = [t for t in ping if t is not m]
        s, b, _ = inf["after_res/conv"]
        m = chk(T("after_res/conv", lambda: ops.conv2d(m, pk["after_res/conv"], f, k, k, out=free[0], scale=s, shift=b, residual=skip)))
        h_, w_, cin = H, W, f
        for i in range(self.upscale_times):
            n, scale = "upscaling/%d/block" % i, 2 ** (i + 1)
            r = scale // 2
            iu = T(n + "/upscaled_concat", lambda: ops.resize_nearest_bilinear_pair(x, r))
            ia = T(n + "/attention", lambda: ops.fold_taps(iu, row_taps(k), out=self._buf("fold_up%d" % i, (B, h_, w_, 64), dt), dtype=dt))
            au = chk(T(n + "/attention", lambda: ops.conv2d(ia, pk[n + "/attention"], cin, k, 1,
                                                            out=self._buf("att_up%d" % i, (B, h_, w_, cin), dt), shift=w[n + "/attention"][1])))
            g = T(n + "/attention_multiply", lambda: ops.gate_mul(au, m, self._buf("gated_up%d" % i, (B, h_, w_, cin), dt)))
            wt, bt = w[n + "/to_add_input_conv_transp"]
            ta = chk(T(n + "/to_add_input_conv_transp", lambda: ops.atanh_conv2d_transpose_small(
                x, wt, bt, scale, out=self._buf("to_add%d" % i, (B, 2 * h_, 2 * w_, ATT_UP_FILTERS), dt))))
            m = chk(T(n + "/conv_transp", lambda: ops.conv2d_transpose_s2(
                g, pk[n + "/conv_transp"], ATT_UP_FILTERS, k, k, out=self._buf("up%d" % i, (B, 2 * h_, 2 * w_, ATT_UP_FILTERS), dt),
                shift=w[n + "/conv_transp"][1], act=ops.ACT_LEAKY, leaky=0.2, residual=ta)))
            h_, w_, cin = 2 * h_, 2 * w_, ATT_UP_FILTERS
        if out is None:
            out = torch.empty((B, h_, w_, 3), dtype=torch.float32, device=self.device)
        return T("final/conv", lambda: ops.conv2d_scatter(m, pk["final/conv"], 3, 9, 9, bias=w["final/conv"][1], act=ops.ACT_TANH, out=out))


class UpscalerIncepResnetNet(UpscalerOrigNet):
    """make_upscaler_incep_resnet (model.py:445-497): a 9x9 head conv WITHOUT activation, 20 inception residual blocks -- every
    path a chain of pre-activation mini blocks BatchNorm -> PReLU -> Conv2D with 1x1, kxk, 1xk and kx1 kernels and 19-64
    filters (model.py:372-443), the paths concatenated and closed by a 1x1 conv -- then the `orig` tail.  Inference path.
    On the B200 kernels every tensor keeps 64 channels (narrower paths are zero-padded: zero weight rows, zero folded scale /
    shift, so padding channels stay exactly 0).  Inside a path the NEXT mini block's BatchNorm + PReLU is the epilogue of the
    conv that feeds it; only the first mini block of a path needs a stand-alone affine + PReLU pass over the block input.
    Concatenate + 1x1 conv is linear in the paths, so it runs as one 1x1 conv per path chained through the epilogue's
    residual operand, the first link carrying the block input (the residual Add): no concatenated tensor exists."""

    supports_training = False
    extra_layers = [("prefinal/tanh", "add"), ("final/tanh", "activation")]

    def __init__(self, filters=64, upscale_factor=4, a=("3path", 5, 3), b=("2path", 10, 7), c=("2path", 5, 3), weights=None,
                 device="cuda", seed=None, infer_dtype=torch.float16):
        self.block_cfg = (tuple(a), tuple(b), tuple(c))
        self.blocks = incep_resnet_blocks(filters, *self.block_cfg)
        super().__init__(kernel_size=c[2], filters=filters, upscale_factor=upscale_factor, res_block_num=len(self.blocks),
                         weights=weights, device=device, seed=seed, infer_dtype=infer_dtype)

    def _layer_specs(self):
        return upscaler_incep_resnet_layer_specs(self.filters, self.upscale_factor, *self.block_cfg)

    def _build_index_maps(self):
        self._final_chan_map, self._convT_wgrad_idx, self._convT_bias_idx = None, {}, {}
        self._final_rows = True

        def padded(kh, kw, cin, cout, c0=0, cin_total=None):
            """[64][kh*kw][64] index map of a Keras kernel (kh, kw, cin_total, cout) restricted to input channels [c0, c0+cin)."""
            ct = cin_total or cin
            idx = -np.ones((64, kh * kw, 64), np.int64)
            t = np.arange(kh * kw)[None, :, None]
            ci = np.arange(cin)[None, None, :]
            co = np.arange(cout)[:, None, None]
            idx[:cout, :, :cin] = (t * ct + c0 + ci) * cout + co
            return idx.astype(np.int32)

        self.add_pack("infer", "initial/conv/9x9", "initial/conv/9x9", ops.conv_rowfold_pack_index(9, 9, 3, self.filters))
        for base, paths, cat in self.blocks:
            c0 = 0
            for pname, minis in paths:
                for name, (kh, kw), cin, cout in minis:
                    lname = name + "/%dx%d" % (kh, kw)
                    self.add_pack("infer", lname, lname, padded(kh, kw, cin, cout))
                width = minis[-1][3]
                self.add_pack("infer", base + "/final/1x1#" + pname, base + "/final/1x1", padded(1, 1, width, self.filters, c0, cat))
                c0 += width
        k = self.kernel_size
        self.add_pack("infer", "prefinal/conv2d", "prefinal/conv2d", ops.conv_pack_index(k, k, 64, 64))
        cin = 64
        for i in range(self.upscale_times):
            n = "upscaling/%d/block/conv_transp" % i
            self.add_pack("infer", n, n, ops.convT_pack_index(k, k, cin, 256))
            cin = 256
        self.add_pack("infer", "final/conv", "final/conv", ops.conv_scatter_pack_index(9, 9, cin, 3))
        self.build_packs({"infer": self.infer_dtype})

    def refresh(self, sets=("infer",)):
        for s in sets:
            self.refresh_packs(s)
            setattr(self, "_%s_dirty" % s, False)
        if "infer" not in sets:
            return
        self._sat_pending = True
        w = self.weights
        nslots = sum(1 + len(m) for _, paths, _ in self.blocks for _, m in paths) + 1     # per path: its pre-activation + its convs
        vec = self._vec("fold", 4 * 64 * nslots, zero=True)                               # [scale | shift | alpha | bias] x 64 per slot
        slot = [0]

        def take():
            i = slot[0]
            slot[0] += 1
            return [vec[4 * 64 * i + 64 * j:4 * 64 * i + 64 * (j + 1)] for j in range(4)]

        def fold(bn_name, conv_bias, scale, shift):
            gamma, beta, mean, var = w[bn_name]
            C = gamma.numel()
            ops.bn_fold(gamma, beta, mean, var, conv_bias, scale[:C], shift[:C], BN_EPS)

        inf = {}
        for base, paths, cat in self.blocks:
            for pname, minis in paths:
                # stand-alone BatchNorm + PReLU of the path's first mini block (applied to the block input)
                sc, sh, al, _ = take()
                fold(minis[0][0] + "/batch_norm", None, sc, sh)
                a0 = w[minis[0][0] + "/prelu"][0].reshape(-1)
                al[:a0.numel()].copy_(a0)
                inf[minis[0][0] + "#pre"] = (sc, sh, al)
                for j, (name, (kh, kw), cin, cout) in enumerate(minis):
                    lname = name + "/%dx%d" % (kh, kw)
                    sc, sh, al, bi = take()
                    if j + 1 < len(minis):          # the conv's epilogue = the next mini block's BatchNorm + PReLU
                        nxt = minis[j + 1][0]
                        fold(nxt + "/batch_norm", w[lname][1], sc, sh)
                        an = w[nxt + "/prelu"][0].reshape(-1)
                        al[:an.numel()].copy_(an)
                        inf[lname] = (sc, sh, al, ops.ACT_PRELU)
                    else:                           # last conv of the path: bias only
                        bi[:cout].copy_(w[lname][1])
                        inf[lname] = (None, bi, None, ops.ACT_NONE)
            inf[base + "/final/1x1"] = w[base + "/final/1x1"][1]
        sc, sh, _, _ = take()
        fold("prefinal/batch_norm", w["prefinal/conv2d"][1], sc, sh)
        inf["prefinal/conv2d"] = (sc, sh)
        for i in range(self.upscale_times):
            n = "upscaling/%d/block/conv_transp" % i
            inf[n] = w[n][1]
        self._infer = inf

    def _sat_names(self):
        return (["initial/conv/9x9"] + [b[0] for b in self.blocks] + ["prefinal/conv2d"] +
                ["upscaling/%d/block/conv_transp" % i for i in range(self.upscale_times)])

    def forward_train(self, *a, **k):
        raise NotImplementedError("make_upscaler_incep_resnet: the inference path is built on the B200 kernels; its training "
                                  "step is not (use make_upscaler_orig or make_upscaler_skip_con for training)")

    backward = forward_train

    def _forward_infer(self, x, out, counters):
        B, H, W, C = x.shape
        f, dt, k = self.filters, self.infer_dtype, self.kernel_size
        inf, pk, w = self._infer, self._packed["infer"], self.weights
        slot = [0]

        def chk(t):
            if counters is not None:
                ops.count_saturated(t, counters[slot[0]:slot[0] + 1])
                slot[0] += 1
            return t
        T = self._timed
        buf = lambda name: self._buf(name, (B, H, W, f), dt)
        xp = T("initial/fold", lambda: ops.fold_taps(x, [(0, dx - 4) for dx in range(9)], out=self._buf("fold", (B, H, W, 64), dt), dtype=dt))
        skip = chk(T("initial/conv", lambda: ops.conv2d(xp, pk["initial/conv/9x9"], f, 9, 1, out=buf("skip"), shift=w["initial/conv/9x9"][1])))
        m, streams = skip, [buf("s0"), buf("s1")]
        for bi, (base, paths, cat) in enumerate(self.blocks):
            acc = m                                     # the chain of 1x1 convs starts from the block input (the residual Add)
            for pi, (pname, minis) in enumerate(paths):
                sc, sh, al = inf[minis[0][0] + "#pre"]
                t = T("inc_res_block/*/pre", lambda: ops.affine_act(m, buf("p0"), sc, sh, al, ops.ACT_PRELU))
                for j, (name, (kh, kw), cin, cout) in enumerate(minis):
                    lname = name + "/%dx%d" % (kh, kw)
                    sc, sh, al, act = inf[lname]
                    t = T("inc_res_block/*/conv", lambda: ops.conv2d(t, pk[lname], f, kh, kw, out=buf("p%d" % (1 + j % 2)), scale=sc, shift=sh,
                                                                     alpha=al, act=act))
                last = pi == len(paths) - 1
                dst = [s_ for s_ in streams if s_ is not m][0] if last else buf("c%d" % (pi % 2))
                acc = T("inc_res_block/*/final_1x1", lambda: ops.conv2d(t, pk[base + "/final/1x1#" + pname], f, 1, 1, out=dst,
                                                                        shift=inf[base + "/final/1x1"] if pi == 0 else None, residual=acc))
            m = chk(acc)
        sc, sh = inf["prefinal/conv2d"]
        nxt = [s_ for s_ in streams if s_ is not m][0]
        m = chk(T("prefinal/conv2d", lambda: ops.conv2d(m, pk["prefinal/conv2d"], 64, k, k, out=nxt, scale=sc, shift=sh, residual=skip)))
        h, w_ = H, W
        for i in range(self.upscale_times):
            n = "upscaling/%d/block/conv_transp" % i
            ob = self._buf("up%d" % i, (B, 2 * h, 2 * w_, 256), dt)
            m = chk(T(n, lambda: ops.conv2d_transpose_s2(m, pk[n], 256, k, k, out=ob, shift=inf[n], act=ops.ACT_LEAKY, leaky=0.2)))
            h, w_ = 2 * h, 2 * w_
        if out is None:
            out = torch.empty((B, h, w_, 3), dtype=torch.float32, device=self.device)
        return T("final/conv", lambda: ops.conv2d_scatter(m, pk["final/conv"], 3, 9, 9, bias=w["final/conv"][1], act=ops.ACT_TANH, out=out))


class Vgg19Net(Net):
    """keras.applications VGG19(include_top=False) up to block5_conv4 (post-ReLU), frozen, fed the raw [-1,1] RGB
    tensor (model.py:108-117).  Forward on either element type; dgrad back to the image for the y_pred branch."""

    input_name = "input_1"
    extra_layers = []
    # The loss compares two feature maps that converge towards each other, so feature rounding noise biases it upwards
    # (measured: bf16 features move a converged loss by up to 2 %, fp16 by < 0.1 %).  The frozen VGG therefore stores
    # activations and packed weights as IEEE fp16; its data gradients are carried multiplied by `grad_scale` (the loss
    # gradient is formed without the 1/n of the mean) so they sit in fp16's normal range, and the scale is divided out in
    # the fp32 epilogue of the last data-gradient convolution.
    train_dtype = torch.float16

    def __init__(self, weights=None, device="cuda", seed=None):
        super().__init__(vgg19_layer_specs(), weights, seed, device)
        for name, kind, c in self.specs:
            if c["cin"] < 64:
                self.add_pack("train", name, name, ops.conv_fullfold_pack_index(3, 3, c["cin"], c["cout"]))
                self.add_pack("train", name + "#dgrad", name, ops.conv_dgrad_pack_index(3, 3, c["cin"], c["cout"]))
            else:
                self.add_pack("train", name, name, ops.conv_pack_index(3, 3, c["cin"], c["cout"]))
                self.add_pack("train", name + "#dgrad", name, ops.conv_dgrad_pack_index(3, 3, c["cin"], c["cout"]))
        self.build_packs({"train": self.train_dtype})
        if self.on_gpu:
            self.refresh()

    def refresh(self, sets=("train",)):
        for s in sets:
            self.refresh_packs(s)

    def output_shape_for(self, input_shape):
        h, w, _ = input_shape
        for _ in range(4):
            h, w = (None if h is None else h // 2), (None if w is None else w // 2)
        return (h, w, 512)

    def forward(self, x, save=False):
        """x: (B,H,W,3) fp32 -> block5_conv4 features (B,H/16,W/16,512) 16-bit; `save` keeps the activations for
        backward()."""
        self._need_gpu()
        dt, pk, w = self.train_dtype, self._packed["train"], self.weights
        B, H, W, _ = x.shape
        taps = [(ky - 1, kx - 1) for ky in range(3) for kx in range(3)]
        a = ops.fold_taps(x, taps, dtype=dt)
        acts, i, first = [], 0, True
        for c in VGG19_CFG:
            if c == "M":
                p = ops.maxpool2x2(a)
                acts.append(("pool", a, p))
                a = p
                continue
            name = VGG19_NAMES[i]
            kk = 1 if first else 3
            a = ops.conv2d(a, pk[name], c, kk, kk, shift=w[name][1], act=ops.ACT_RELU)
            acts.append(("conv", name, a))
            first, i = False, i + 1
        return (a, acts) if save else a

    def forward_infer(self, x, out=None):
        return self.forward(x)

    def backward(self, acts, dfeat, grad_scale=1.0):
        """dfeat: grad_scale x the gradient w.r.t. the (post-ReLU) block5_conv4 features, 16-bit.  Returns the gradient
        w.r.t. the input image, fp32 (B,H,W,3), with the scale divided out.  Weights are frozen: data gradients only."""
        pk = self._packed["train"]
        top = acts[-1][2]
        # ReLU backward of the top layer
        d = ops.bn_bwd_apply(top, dfeat, None, None, None, ops.ACT_RELU, 0.0, None, None, None, None, 0, torch.empty_like(top))
        for j in range(len(acts) - 1, 0, -1):
            kind, name, _ = acts[j]
            below = acts[j - 1]
            if kind == "pool":
                continue        # handled together with the conv above it
            cin = pk[name + "#dgrad"].shape[0]
            if below[0] == "conv":
                d = ops.conv2d(d, pk[name + "#dgrad"], cin, 3, 3, gate=below[2], gate_slope=0.0)
            else:
                _, a_in, p_out = below
                dp = ops.conv2d(d, pk[name + "#dgrad"], cin, 3, 3)
                d = ops.maxpool2x2_bwd(a_in, p_out, dp, relu_gate=True)
        name = acts[0][1]
        B, H, W, _ = d.shape
        inv = self._vec("inv_scale", 3)
        inv.fill_(1.0 / grad_scale)
        return ops.conv2d(d, pk[name + "#dgrad"], 3, 3, 3, scale=inv,
                          out=torch.empty((B, H, W, 3), dtype=torch.float32, device=self.device))


class DiscriminatorSimple512Net(Net):
    """make_discriminator_simple_512 (model.py:836-896): 9 x (3x3 conv [s1 then 8 x s2] + BN + LeakyReLU 0.1) ->
    Flatten -> Dense 1024 - BN - LReLU -> Dense 32 - BN - LReLU -> Dense 1 -> optional activation."""

    input_name = "discriminator/input"
    filters = DISC_FILTERS

    def __init__(self, input_shape, activation="none", weights=None, device="cuda", seed=None, dtype=None):
        self.input_shape = tuple(input_shape)
        self.activation = activation
        self.train_dtype = dtype or TRAIN_DTYPE
        specs = self._layer_specs()
        super().__init__(specs, weights, seed, device)
        self.extra_layers = [("discriminator/final/Flatten", "flatten")]
        self._setup_convs(specs)
        self._dense = [(n, c) for n, kind, c in specs if kind == "dense"]
        self.build_packs({"train": self.train_dtype})
        if self.on_gpu:
            self.refresh()

    def _layer_specs(self):
        return discriminator_simple_512_layer_specs(self.input_shape, self.filters)

    def _setup_convs(self, specs):
        h, w = self.input_shape[:2]
        self._conv_geom = []
        for name, kind, c in specs:
            if kind != "conv":
                continue
            s = c["stride"]
            oh, ow = -(-h // s), -(-w // s)
            self._conv_geom.append((name, c, (h, w), (oh, ow)))
            h, w = oh, ow
            if c["cin"] < 64:
                self.add_pack("train", name, name, ops.conv_fullfold_pack_index(3, 3, c["cin"], c["cout"]))
                self.add_pack("train", name + "#dgrad", name, ops.conv_dgrad_pack_index(3, 3, c["cin"], c["cout"]))
            else:
                self.add_pack("train", name, name, ops.conv_pack_index(3, 3, c["cin"], c["cout"]))
                self.add_pack("train", name + "#dgrad", name, ops.conv_s2_dgrad_pack_index(3, 3, c["cin"], c["cout"]))
        self._flat_hw = (h, w)
        c1 = specs[0][2]
        self._first_wgrad_idx = torch.from_numpy(ops.fullfold_wgrad_gather_index(3, 3, c1["cin"], c1["cout"])).to(self.device)

    def refresh(self, sets=("train",)):
        for s in sets:
            self.refresh_packs(s)

    def output_shape_for(self, input_shape):
        return (1,)

    def _bn_act(self, tape, bn_name, z, training, sync, update_moving=True):
        """BatchNormalization + LeakyReLU(0.1): batch statistics when training, moving statistics otherwise."""
        out = torch.empty_like(z)
        if training:
            return self._bn_forward(tape, bn_name, bn_name, z, out, ops.ACT_LEAKY, leaky=0.1, sync=sync,
                                    update_moving=update_moving)
        C = z.shape[-1]
        aff = self._vec("aff/" + bn_name, 4 * C)
        g, b, mm, mv = self.weights[bn_name]
        ops.bn_fold(g, b, mm, mv, None, aff[:C], aff[C:2 * C], BN_EPS)
        return ops.affine_act(z, out, aff[:C], aff[C:2 * C], None, ops.ACT_LEAKY, 0.1)

    def forward(self, x, training, sync=None, update_moving=True):
        """x: (n,H,W,3) fp32 -> ((n,1) fp32 scores, tape).  training=True normalises with batch statistics and, when
        `update_moving`, moves the moving ones (training.py decides that per compiled model); False = Model.predict."""
        self._need_gpu()
        dt, pk, w = self.train_dtype, self._packed["train"], self.weights
        n = x.shape[0]
        if tuple(x.shape[1:]) != self.input_shape:
            raise ValueError("discriminator expects input %s, got %s" % (self.input_shape, tuple(x.shape[1:])))
        tape = {}
        a = self._conv_forward(x, tape, training, sync, update_moving)
        a = a.reshape(n, -1)                                   # Keras Flatten of an NHWC tensor
        dins = []
        for j, (name, c) in enumerate(self._dense):
            dins.append(a)
            last = j == len(self._dense) - 1
            z = ops.dense_fwd(a, w[name][0], w[name][1],
                              torch.empty((n, c["cout"]), dtype=torch.float32 if last else dt, device=self.device))
            a = z if last else self._bn_act(tape, "discriminator/final/BatchNorm_%d" % (j + 1), z, training, sync, update_moving)
        tape["dins"] = dins
        if self.activation in ops.SCORE_ACTS:                   # model.py:885-892; any other string: no output layer
            tape["score_pre"] = a
            a = ops.score_act(a, self.activation)
        return a, tape

    def _conv_forward(self, x, tape, training, sync, update_moving=True):
        """The convolutional blocks: block_1 with all nine taps folded into channels, then 3x3 'same' convs."""
        dt, pk, w = self.train_dtype, self._packed["train"], self.weights
        taps = [(ky - 1, kx - 1) for ky in range(3) for kx in range(3)]
        a = ops.fold_taps(x, taps, dtype=dt)
        ins = []
        for bi, (name, c, _, _) in enumerate(self._conv_geom):
            kk = 1 if bi == 0 else 3
            ins.append(a)
            z = ops.conv2d(a, pk[name], c["cout"], kk, kk, stride=c["stride"], shift=w[name][1])
            a = self._bn_act(tape, name.replace("Conv2d", "BatchNorm"), z, training, sync, update_moving)
        tape["ins"] = ins
        return a

    def forward_infer(self, x, out=None):
        return self.forward(x, False)[0]

    def backward(self, tape, dscore, want_wgrad=True, want_input_grad=False, sync=None, on_ready=None, grad_scale=1.0):
        """dscore: (n,1) fp32 gradient w.r.t. the scores.  Accumulates weight gradients unless the discriminator is
        frozen (gan_train, model.py:1040); returns the gradient w.r.t. the input image (fp32) when asked.
        `on_ready(offset)`: see UpscalerOrigNet.backward (the dense head, most of the parameters, is ready first).
        `grad_scale`: dscore carries this factor on top of the step's scale (fp16 range inside a frozen discriminator whose
        adversarial weight is tiny, train_gan.py -dw 1e-10); it is divided out of the returned image gradient in fp32."""
        by_tensor = torch.is_tensor(grad_scale)      # a 64-float device vector holding 1 / scale: the caller fills it before
        if want_wgrad and (by_tensor or grad_scale != 1.0):     # each replay of a captured graph (training._step_gan2_graph)
            raise ValueError("an internal gradient scale is only supported for a frozen discriminator")
        self._inv_gscale = None
        if by_tensor:
            self._inv_gscale = grad_scale
        elif grad_scale != 1.0:
            self._inv_gscale = self._vec("inv_gscale", 64)
            self._inv_gscale.fill_(1.0 / grad_scale)
        ready = on_ready if (on_ready is not None and want_wgrad) else (lambda off: None)
        dt, pk, w = self.train_dtype, self._packed["train"], self.weights
        if want_wgrad:
            self.ensure_grads()
        g = self.grads
        n = dscore.shape[0]
        dins = tape["dins"]
        d = dscore
        if "score_pre" in tape:
            d = ops.score_act(tape["score_pre"], self.activation, dy=dscore.contiguous())
        for j in reversed(range(len(self._dense))):
            name, c = self._dense[j]
            if j < len(self._dense) - 1:
                bn = "discriminator/final/BatchNorm_%d" % (j + 1)
                d = self._bn_backward(tape, bn, bn, d, torch.empty_like(d), ops.ACT_LEAKY, leaky=0.1, sync=sync,
                                      want_param_grads=want_wgrad)
            if want_wgrad:
                ops.dense_bwd_weight(dins[j], d, g[name][0], g[name][1])
            d = ops.dense_bwd_data(d, w[name][0], torch.empty((n, c["cin"]), dtype=dt, device=self.device))
        ready(self.offsets[(self._dense[0][0], 0)][0])
        self.probe_grad(d)
        last = self._conv_geom[-1]
        d = d.view(n, last[3][0], last[3][1], last[1]["cout"])
        return self._conv_backward(tape, d, want_wgrad, want_input_grad, sync, ready)

    def _conv_backward(self, tape, d, want_wgrad, want_input_grad, sync, ready):
        dt, pk, w = self.train_dtype, self._packed["train"], self.weights
        g = self.grads
        n = d.shape[0]
        ins = tape["ins"]
        for bi in reversed(range(len(self._conv_geom))):
            name, c, (h, w_), _ = self._conv_geom[bi]
            bn = name.replace("Conv2d", "BatchNorm")
            dz = self._bn_backward(tape, bn, bn, d, torch.empty_like(d), ops.ACT_LEAKY, leaky=0.1, sync=sync,
                                   want_param_grads=want_wgrad)
            if bi % 2 == 1:
                self.probe_grad(dz)
            if bi > 0:
                if want_wgrad:
                    ops.conv2d_wgrad(ins[bi], dz, 3, 3, c["stride"], g[name][0])
                d = ops.conv2d_s2_dgrad(dz, pk[name + "#dgrad"], c["cin"], 3, 3, (h, w_))
            else:
                if want_wgrad:
                    raw = self._buf("first_raw", (64 * c["cout"],), torch.float32)
                    raw.zero_()
                    ops.conv2d_wgrad(ins[0], dz, 1, 1, 1, raw)
                    ops.gather_f32(raw, self._first_wgrad_idx, g[name][0], accumulate=True)
                ready(0)
                if want_input_grad:
                    return ops.conv2d(dz, pk[name + "#dgrad"], c["cin"], 3, 3, scale=self._inv_gscale,
                                      out=torch.empty((n, h, w_, c["cin"]), dtype=torch.float32, device=self.device))
        return None


class DiscriminatorThin512Net(DiscriminatorSimple512Net):
    """make_discriminator_thin_512 (model.py:901-960): the same graph with 128-wide blocks 2..9."""

    filters = DISC_THIN_FILTERS


class DiscriminatorSparse512Net(DiscriminatorSimple512Net):
    """make_discriminator_sparse_512 (model.py:964-1012): six 5x5 padding='valid' blocks with strides 1,3,3,3,3,3 and a
    Dense 128 / 32 / 1 head.  block_1 (Cin = 3) runs as a 5x1 conv over the input with its 5 filter columns folded into
    channels; the stride-3 blocks run on b2_conv2d_fprop_pad / b2_conv2d_strided_dgrad / b2_conv2d_wgrad_pad."""

    filters = DISC_SPARSE_FILTERS

    def _layer_specs(self):
        return discriminator_sparse_512_layer_specs(self.input_shape)

    def _setup_convs(self, specs):
        h, w = self.input_shape[:2]
        self._conv_geom = []
        for name, kind, c in specs:
            if kind != "conv":
                continue
            s = c["stride"]
            oh, ow = (h - 5) // s + 1, (w - 5) // s + 1
            self._conv_geom.append((name, c, (h, w), (oh, ow)))
            h, w = oh, ow
            if c["cin"] < 64:
                self.add_pack("train", name, name, ops.conv_rowfold_pack_index(5, 5, c["cin"], c["cout"]))
                self.add_pack("train", name + "#dgrad", name, ops.conv_dgrad_pack_index(5, 5, c["cin"], c["cout"]))
            else:
                self.add_pack("train", name, name, ops.conv_pack_index(5, 5, c["cin"], c["cout"]))
                self.add_pack("train", name + "#dgrad", name, ops.conv_s2_dgrad_pack_index(5, 5, c["cin"], c["cout"]))
        self._flat_hw = (h, w)
        c1 = specs[0][2]
        self._first_wgrad_idx = torch.from_numpy(ops.rowfold_wgrad_gather_index(5, 5, c1["cin"], c1["cout"])).to(self.device)

    def _conv_forward(self, x, tape, training, sync, update_moving=True):
        dt, pk, w = self.train_dtype, self._packed["train"], self.weights
        H, W = self.input_shape[:2]
        a = ops.fold_taps(x, [(0, kx) for kx in range(5)], out_hw=(H, W - 4), dtype=dt)
        ins = []
        for bi, (name, c, _, ohw) in enumerate(self._conv_geom):
            ins.append(a)
            z = ops.conv2d_pad(a, pk[name], c["cout"], 5, 1 if bi == 0 else 5, c["stride"], (0, 0), ohw, shift=w[name][1])
            a = self._bn_act(tape, name.replace("Conv2d", "BatchNorm"), z, training, sync, update_moving)
        tape["ins"] = ins
        return a

    def _conv_backward(self, tape, d, want_wgrad, want_input_grad, sync, ready):
        pk, g = self._packed["train"], self.grads
        n = d.shape[0]
        ins = tape["ins"]
        for bi in reversed(range(len(self._conv_geom))):
            name, c, (h, w_), _ = self._conv_geom[bi]
            bn = name.replace("Conv2d", "BatchNorm")
            dz = self._bn_backward(tape, bn, bn, d, torch.empty_like(d), ops.ACT_LEAKY, leaky=0.1, sync=sync,
                                   want_param_grads=want_wgrad)
            if bi % 2 == 1:
                self.probe_grad(dz)
            if bi > 0:
                if want_wgrad:
                    ops.conv2d_wgrad_pad(ins[bi], dz, 5, 5, c["stride"], (0, 0), g[name][0])
                d = ops.conv2d_strided_dgrad(dz, pk[name + "#dgrad"], c["cin"], 5, 5, c["stride"], (0, 0), (h, w_))
            else:
                if want_wgrad:
                    raw = self._buf("first_raw", (5 * 64 * c["cout"],), torch.float32)
                    raw.zero_()
                    ops.conv2d_wgrad_pad(ins[0], dz, 5, 1, 1, (0, 0), raw)
                    ops.gather_f32(raw, self._first_wgrad_idx, g[name][0], accumulate=True)
                ready(0)
                if want_input_grad:
                    return ops.conv2d_pad(dz, pk[name + "#dgrad"], c["cin"], 5, 5, 1, (4, 4), (h, w_), scale=self._inv_gscale,
                                          out=torch.empty((n, h, w_, c["cin"]), dtype=torch.float32, device=self.device))
        return None
