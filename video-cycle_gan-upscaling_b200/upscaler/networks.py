"""Device-side networks of the hot path, executed layer by layer through the C ABI.

UpscalerOrigNet is the graph make_upscaler_orig declares (reference: upscaling/upscaler/model.py:267-295, with
residual_block :15-27 and upsampling_block :70-75).  Weights are kept as fp32 master tensors in Keras layout under
the reference's layer names (so .h5 files map one-to-one) plus packed bf16 copies in the kernels' layout.
"""
import math
from collections import OrderedDict

import numpy as np
import torch

from . import ops

BN_EPS = 1e-3  # keras BatchNormalization default, model.py:20


def glorot_uniform(rng, shape, fan_in, fan_out):
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def upscaler_orig_layer_specs(kernel_size=5, filters=64, upscale_factor=4, res_block_num=16):
    """(name, kind, config) in graph order, with the reference's layer names."""
    k = kernel_size
    specs = [("initial/conv", "conv", dict(kh=9, kw=9, cin=3, cout=filters)),
             ("initial/prelu", "prelu", dict(c=filters))]
    for i in range(res_block_num):
        n = "res_block/%d" % i
        specs += [(n + "/conv_pre", "conv", dict(kh=k, kw=k, cin=filters, cout=filters)),
                  (n + "/batch_norm_pre", "bn", dict(c=filters)),
                  (n + "/prelu", "prelu", dict(c=filters)),
                  (n + "/conv_post", "conv", dict(kh=k, kw=k, cin=filters, cout=filters)),
                  (n + "/batch_norm_post", "bn", dict(c=filters))]
    specs += [("prefinal/conv2d", "conv", dict(kh=k, kw=k, cin=filters, cout=64)),
              ("prefinal/batch_norm", "bn", dict(c=64))]
    cin = 64
    for i in range(int(math.log(upscale_factor, 2))):
        specs.append(("upscaling/%d/block/conv_transp" % i, "convT", dict(kh=k, kw=k, cin=cin, cout=256)))
        cin = 256
    specs.append(("final/conv", "conv", dict(kh=9, kw=9, cin=cin, cout=3)))
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


class UpscalerOrigNet:
    """SRGAN generator: 9x9 conv + PReLU -> N x (conv-BN-PReLU-conv-BN-add) -> conv-BN-add(skip) ->
    log2(factor) x (ConvT s2 + LeakyReLU 0.2) -> 9x9 conv -> tanh."""

    def __init__(self, kernel_size=5, filters=64, upscale_factor=4, res_block_num=16, weights=None, device="cuda",
                 seed=None, infer_dtype=torch.float16):
        if filters != 64:
            raise ValueError("the B200 path is built for filters=64 (reference default, model.py:267)")
        self.kernel_size, self.filters = kernel_size, filters
        self.upscale_factor, self.res_block_num = upscale_factor, res_block_num
        self.upscale_times = int(math.log(upscale_factor, 2))
        # Without a CUDA device the object still carries structure and weights (shapes, parameter counts, .h5 I/O);
        # anything that computes raises, there is no CPU execution path.
        self.on_gpu = torch.cuda.is_available()
        self.device = torch.device(device) if self.on_gpu else torch.device("cpu")
        # 16-bit storage type of the inference path (activations + packed weights): fp16 keeps 11 significand bits,
        # which is what holds max-abs error under 2e-2 across 37 layers; bf16 is accepted and ~8x noisier.
        self.infer_dtype = infer_dtype
        self.specs = upscaler_orig_layer_specs(kernel_size, filters, upscale_factor, res_block_num)
        host = weights if weights is not None else init_keras_weights(self.specs, seed)
        self.weights = OrderedDict()
        for name, kind, cfg in self.specs:
            if name not in host:
                raise KeyError("missing weights for layer %s" % name)
            self.weights[name] = [torch.as_tensor(np.asarray(a), dtype=torch.float32).to(self.device).contiguous()
                                  for a in host[name]]
        self._pack_idx = {}
        self._packed = {}
        self._infer = {}
        self._buffers = {}
        self._final_rows = False
        self.layer_events = None  # dict name -> [(start, end) CUDA events] when bench.py enables per-layer timing
        self._build_index_maps()
        if self.on_gpu:
            self.refresh()

    # ---------------------------------------------------------------- weights
    def _build_index_maps(self):
        for name, kind, c in self.specs:
            if kind == "conv":
                if name == "final/conv" and c["kh"] == 9 and c["cout"] == 3 and c["cin"] % 64 == 0 and c["cin"] <= 256:
                    idx = ops.conv_rows_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
                    self._final_rows = True
                elif c["cin"] < 64:
                    idx = ops.conv_rowfold_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
                else:
                    idx = ops.conv_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
            elif kind == "convT":
                idx = ops.convT_pack_index(c["kh"], c["kw"], c["cin"], c["cout"])
            else:
                continue
            self._pack_idx[name] = torch.from_numpy(idx).to(self.device)

    def refresh(self):
        """Re-derive the packed bf16 weights and the folded inference epilogues from the fp32 masters."""
        for name, idx in self._pack_idx.items():
            self._packed[name] = ops.gather_cast(self.weights[name][0], idx, self._packed.get(name), self.infer_dtype)
        w = self.weights
        inf = {}

        def fold_bn(conv, bn):
            gamma, beta, mean, var = w[bn]
            scale = gamma / torch.sqrt(var + BN_EPS)
            shift = beta + (w[conv][1] - mean) * scale
            return scale.contiguous(), shift.contiguous()

        inf["initial/conv"] = (None, w["initial/conv"][1], w["initial/prelu"][0].reshape(-1).contiguous())
        for i in range(self.res_block_num):
            n = "res_block/%d" % i
            s, b = fold_bn(n + "/conv_pre", n + "/batch_norm_pre")
            inf[n + "/conv_pre"] = (s, b, w[n + "/prelu"][0].reshape(-1).contiguous())
            s, b = fold_bn(n + "/conv_post", n + "/batch_norm_post")
            inf[n + "/conv_post"] = (s, b, None)
        s, b = fold_bn("prefinal/conv2d", "prefinal/batch_norm")
        inf["prefinal/conv2d"] = (s, b, None)
        for i in range(self.upscale_times):
            n = "upscaling/%d/block/conv_transp" % i
            inf[n] = (None, w[n][1], None)
        inf["final/conv"] = (None, w["final/conv"][1], None)
        self._infer = inf

    input_name = "initial/input"
    extra_layers = [("prefinal/tanh", "add"), ("final/tanh", "activation")]

    def output_shape_for(self, input_shape):
        h, w, c = input_shape
        f = 2 ** self.upscale_times
        return (None if h is None else h * f, None if w is None else w * f, 3)

    def host_weights(self):
        return OrderedDict((k, [t.detach().cpu().numpy() for t in ts]) for k, ts in self.weights.items())

    def count_params(self):
        return int(sum(t.numel() for ts in self.weights.values() for t in ts))

    # ---------------------------------------------------------------- inference
    def _buf(self, key, shape, dtype):
        b = self._buffers.get(key)
        if b is None or tuple(b.shape) != tuple(shape) or b.dtype != dtype:
            b = torch.empty(shape, dtype=dtype, device=self.device)
            self._buffers[key] = b
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

    def forward_infer(self, x, out=None):
        """x: (B,H,W,3) float32 CUDA in [-1,1] -> (B, H*f, W*f, 3) float32.  BatchNorm uses moving statistics folded
        into the conv epilogues (Keras learning_phase=0, i.e. Model.predict)."""
        if not self.on_gpu:
            raise RuntimeError("the generator runs on hand-written sm_100a kernels only; no CUDA device is available")
        B, H, W, C = x.shape
        if C != 3 or x.dtype != torch.float32:
            raise ValueError("expected (B,H,W,3) float32 input, got %s %s" % (tuple(x.shape), x.dtype))
        k, f, dt = self.kernel_size, self.filters, self.infer_dtype
        inf, pk = self._infer, self._packed
        # head: fold the 9 horizontal taps of the 3-channel input into 27 (of 64) channels, then a 9x1 conv
        T = self._timed
        xp = T("initial/fold", lambda: ops.fold_taps(x, [(0, dx - 4) for dx in range(9)],
                                                     out=self._buf("fold", (B, H, W, 64), dt), dtype=dt))
        s, b, a = inf["initial/conv"]
        skip = T("initial/conv", lambda: ops.conv2d(xp, pk["initial/conv"], f, 9, 1,
                                                    out=self._buf("skip", (B, H, W, f), dt), scale=s, shift=b, alpha=a,
                                                    act=ops.ACT_PRELU))
        m = skip
        ping = [self._buf("t%d" % i, (B, H, W, f), dt) for i in range(3)]
        for i in range(self.res_block_num):
            n = "res_block/%d" % i
            free = [t for t in ping if t is not m]
            s, b, a = inf[n + "/conv_pre"]
            h = T("res_block/*/conv_pre", lambda: ops.conv2d(m, pk[n + "/conv_pre"], f, k, k, out=free[0], scale=s, shift=b,
                                                             alpha=a, act=ops.ACT_PRELU))
            s, b, _ = inf[n + "/conv_post"]
            m = T("res_block/*/conv_post", lambda: ops.conv2d(h, pk[n + "/conv_post"], f, k, k, out=free[1], scale=s,
                                                              shift=b, residual=m))
        free = [t for t in ping if t is not m]
        s, b, _ = inf["prefinal/conv2d"]
        m = T("prefinal/conv2d", lambda: ops.conv2d(m, pk["prefinal/conv2d"], 64, k, k, out=free[0], scale=s, shift=b,
                                                    residual=skip))
        h, w_ = H, W
        for i in range(self.upscale_times):
            n = "upscaling/%d/block/conv_transp" % i
            _, b, _ = inf[n]
            ob = self._buf("up%d" % i, (B, 2 * h, 2 * w_, 256), dt)
            m = T(n, lambda: ops.conv2d_transpose_s2(m, pk[n], 256, k, k, out=ob, shift=b, act=ops.ACT_LEAKY, leaky=0.2))
            h, w_ = 2 * h, 2 * w_
        _, b, _ = inf["final/conv"]
        if out is None:
            out = torch.empty((B, h, w_, 3), dtype=torch.float32, device=self.device)
        if self._final_rows:
            return T("final/conv", lambda: ops.conv2d_rows(m, pk["final/conv"], 3, 9, 9, bias=b, act=ops.ACT_TANH, out=out))
        return T("final/conv", lambda: ops.conv2d(m, pk["final/conv"], 3, 9, 9, out=out, shift=b, act=ops.ACT_TANH))
