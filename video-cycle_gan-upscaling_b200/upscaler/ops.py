"""Tensor-level wrappers over the C ABI: validate geometry, allocate outputs, call libb2srgan.so.

Tensors are torch CUDA tensors used purely as device-memory handles (NHWC, contiguous).  Every function here ends in
a hand-written sm_100a kernel; nothing falls back to torch ops.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import ACT_LEAKY, ACT_NONE, ACT_PRELU, ACT_RELU, ACT_TANH  # noqa: F401


_launches = 0  # kernels launched through this module (bench.py reports it as gpu_launches)


def reset_launch_count():
    global _launches
    _launches = 0


def launch_count():
    return _launches


def set_pair_kernels(enable):
    """A/B switch for the CTA-pair (cta_group::2) kernels; returns the previous setting."""
    return bool(_lib.load().b2_set_pair_kernels(1 if enable else 0))


def set_pdl(enable):
    """Programmatic dependent launch of every kernel on / off (include/b2srgan.h b2_set_pdl); returns the old setting."""
    return bool(_lib.load().b2_set_pdl(1 if enable else 0))


def _count(n):
    global _launches
    _launches += n


def _require_cuda(*ts):
    for t in ts:
        if t is not None and (not t.is_cuda or not t.is_contiguous()):
            raise ValueError("expected contiguous CUDA tensors")


def _elem_flag(x, w):
    """16-bit element type of a conv's operands: 0 = bf16, 1 = fp16; activations and weights must agree."""
    if x.dtype != w.dtype or x.dtype not in (torch.bfloat16, torch.float16):
        raise ValueError("expected matching bf16 or fp16 activations and packed weights, got %s / %s" % (x.dtype, w.dtype))
    return 1 if x.dtype == torch.float16 else 0


def tf_same_out(size, stride):
    return -(-size // stride)


def conv2d(x, w_packed, cout, kh, kw, stride=1, out=None, out_dtype=None, scale=None, shift=None, alpha=None,
           act=ACT_NONE, leaky=0.0, residual=None, direct=False, gate=None, gate_slope=0.0, stats_part=None):
    """Conv2D(padding='same') forward (reference: keras Conv2D at model.py:19, 22, 275, 283, 290, 839-871).
    `stats_part` (float32 (bn_part_rows(), 2*cout)): receives the partial per-channel [sum | sum of squares] rows of the
    output for the train-mode BatchNorm that follows (fused into the epilogue where the kernel supports it); finalize with
    nparts = bn_part_rows()."""
    B, H, W, Cin = x.shape
    OH, OW = tf_same_out(H, stride), tf_same_out(W, stride)
    f16 = _elem_flag(x, w_packed)
    if Cin % 64 != 0:
        raise ValueError("conv2d: Cin=%d must be a multiple of 64 (fold narrow inputs first)" % Cin)
    if w_packed.numel() != cout * kh * kw * Cin:
        raise ValueError("conv2d: packed weight has %d elements, expected %d" % (w_packed.numel(), cout * kh * kw * Cin))
    if out is None:
        out = torch.empty((B, OH, OW, cout), dtype=out_dtype or x.dtype, device=x.device)
    elif tuple(out.shape) != (B, OH, OW, cout):
        raise ValueError("conv2d: bad output shape %s" % (tuple(out.shape),))
    if out.dtype not in (x.dtype, torch.float32):
        raise ValueError("conv2d: output must be float32 or the activation dtype")
    if residual is not None and (tuple(residual.shape) != tuple(out.shape) or residual.dtype != x.dtype):
        raise ValueError("conv2d: residual must have the activation dtype and the output's shape")
    if gate is not None and (tuple(gate.shape) != tuple(out.shape) or gate.dtype != x.dtype):
        raise ValueError("conv2d: gate must have the activation dtype and the output's shape")
    _require_cuda(x, w_packed, out, scale, shift, alpha, residual, gate, stats_part)
    if stats_part is not None and (stats_part.dtype != torch.float32 or tuple(stats_part.shape) != (bn_part_rows(), 2 * cout)):
        raise ValueError("conv2d: stats_part must be float32 (%d, %d)" % (bn_part_rows(), 2 * cout))
    ep = _lib.make_epilogue(out, scale, shift, alpha, act, leaky, residual, f16, gate, gate_slope, stats_part)
    lib = _lib.load()
    if direct:
        rc = lib.b2_conv2d_direct(_lib.ptr(x), B, H, W, Cin, _lib.ptr(w_packed), cout, kh, kw, stride, 0,
                                  ctypes.byref(ep), _lib.stream_ptr())
    else:
        rc = lib.b2_conv2d_fprop(_lib.ptr(x), B, H, W, Cin, _lib.ptr(w_packed), cout, kh, kw, stride, ctypes.byref(ep),
                                 _lib.stream_ptr())
    _lib.check(rc, "b2_conv2d_fprop")
    _count(1)
    return out


def new_stats_part(cout, device):
    """Scratch for conv2d(..., stats_part=...) -> (tensor, nparts)."""
    return torch.empty((bn_part_rows(), 2 * cout), dtype=torch.float32, device=device), bn_part_rows()


def conv2d_rows(x, w_rows, cout, kh, kw, bias=None, act=ACT_TANH, out=None):
    """Final Conv2D(3, 9xKW, 'same') + bias + tanh with fp32 output (reference: model.py:290-291), row-streaming
    kernel.  `w_rows` is packed by conv_rows_pack_index."""
    B, H, W, Cin = x.shape
    f16 = _elem_flag(x, w_rows)
    if w_rows.numel() != 32 * kw * Cin:
        raise ValueError("conv2d_rows: packed weight size mismatch")
    if out is None:
        out = torch.empty((B, H, W, cout), dtype=torch.float32, device=x.device)
    elif tuple(out.shape) != (B, H, W, cout) or out.dtype != torch.float32:
        raise ValueError("conv2d_rows: output must be float32 (B,H,W,%d)" % cout)
    _require_cuda(x, w_rows, out, bias)
    rc = _lib.load().b2_conv2d_rows_fprop(_lib.ptr(x), B, H, W, Cin, _lib.ptr(w_rows), cout, kh, kw, _lib.ptr(bias), act,
                                          f16, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "b2_conv2d_rows_fprop")
    _count(1)
    return out


def conv2d_scatter(x, w_taps, cout, kh, kw, bias=None, act=ACT_TANH, out=None):
    """Final Conv2D(3, 9x9, 'same') + bias + tanh with fp32 output (reference: model.py:290-291), scatter-form kernel
    (all taps in the GEMM N dimension).  `w_taps` is packed by conv_scatter_pack_index."""
    B, H, W, Cin = x.shape
    f16 = _elem_flag(x, w_taps)
    if w_taps.numel() != 256 * Cin:
        raise ValueError("conv2d_scatter: packed weight size mismatch")
    if out is None:
        out = torch.empty((B, H, W, cout), dtype=torch.float32, device=x.device)
    elif tuple(out.shape) != (B, H, W, cout) or out.dtype != torch.float32:
        raise ValueError("conv2d_scatter: output must be float32 (B,H,W,%d)" % cout)
    _require_cuda(x, w_taps, out, bias)
    rc = _lib.load().b2_conv2d_scatter_fprop(_lib.ptr(x), B, H, W, Cin, _lib.ptr(w_taps), cout, kh, kw, _lib.ptr(bias), act,
                                             f16, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "b2_conv2d_scatter_fprop")
    _count(1)
    return out


def conv2d_transpose_s2(x, w_packed, cout, kh, kw, out=None, scale=None, shift=None, alpha=None, act=ACT_NONE,
                        leaky=0.0, direct=False, residual=None):
    """Conv2DTranspose(strides=2, padding='same') forward (reference: model.py:72).  `residual` (same shape / type as the
    output) is added after the activation (Add of model.py:97)."""
    B, H, W, Cin = x.shape
    f16 = _elem_flag(x, w_packed)
    if Cin % 64 != 0:
        raise ValueError("conv2d_transpose_s2: Cin=%d must be a multiple of 64" % Cin)
    if w_packed.numel() != cout * kh * kw * Cin:
        raise ValueError("conv2d_transpose_s2: packed weight size mismatch")
    if out is None:
        out = torch.empty((B, 2 * H, 2 * W, cout), dtype=x.dtype, device=x.device)
    elif tuple(out.shape) != (B, 2 * H, 2 * W, cout) or out.dtype != x.dtype:
        raise ValueError("conv2d_transpose_s2: bad output shape / dtype")
    if residual is not None and (tuple(residual.shape) != tuple(out.shape) or residual.dtype != out.dtype):
        raise ValueError("conv2d_transpose_s2: residual must match the output")
    _require_cuda(x, w_packed, out, scale, shift, alpha, residual)
    ep = _lib.make_epilogue(out, scale, shift, alpha, act, leaky, residual, f16)
    lib = _lib.load()
    if direct:
        rc = lib.b2_conv2d_direct(_lib.ptr(x), B, H, W, Cin, _lib.ptr(w_packed), cout, kh, kw, 1, 1, ctypes.byref(ep),
                                  _lib.stream_ptr())
    else:
        rc = lib.b2_conv2d_transpose_s2_fprop(_lib.ptr(x), B, H, W, Cin, _lib.ptr(w_packed), cout, kh, kw,
                                              ctypes.byref(ep), _lib.stream_ptr())
    _lib.check(rc, "b2_conv2d_transpose_s2_fprop")
    _count(4)
    return out


def fold_taps(x, taps, stride=1, out_hw=None, out=None, dtype=torch.bfloat16, out_c0=0):
    """out[b,y,x, out_c0 + t*Cin + c] = x[b, y*stride + dy_t, x*stride + dx_t, c]; 64 channels written (zero padded) at
    channel offset out_c0 of `out` (which may be wider than 64 channels)."""
    B, H, W, Cin = x.shape
    if x.dtype not in (torch.float32, torch.bfloat16, torch.float16):
        raise ValueError("fold_taps expects float32, bf16 or fp16 input")
    OH, OW = out_hw if out_hw is not None else (tf_same_out(H, stride), tf_same_out(W, stride))
    if out is None:
        out = torch.empty((B, OH, OW, 64), dtype=dtype, device=x.device)
    if x.dtype != torch.float32 and x.dtype != out.dtype:
        raise ValueError("fold_taps: 16-bit input and output types must match")
    _require_cuda(x, out)
    n = len(taps)
    dy = (ctypes.c_int * n)(*[int(t[0]) for t in taps])
    dx = (ctypes.c_int * n)(*[int(t[1]) for t in taps])
    rc = _lib.load().b2_fold_taps(_lib.ptr(x), 1 if x.dtype == torch.float32 else 0, B, H, W, Cin, OH, OW, stride, n,
                                  dy, dx, _lib.ptr(out), 1 if out.dtype == torch.float16 else 0, out.shape[-1], out_c0,
                                  _lib.stream_ptr())
    _lib.check(rc, "b2_fold_taps")
    _count(1)
    return out


def u8_to_f32_norm(x, out=None):
    """uint8 frames -> float32 in [-1, 1]: x / 127.5 - 1 (data.py:266-270), bit-exact with NumPy's float64 evaluation."""
    if x.dtype != torch.uint8:
        raise ValueError("u8_to_f32_norm expects uint8 frames")
    if out is None:
        out = torch.empty(x.shape, dtype=torch.float32, device=x.device)
    _require_cuda(x, out)
    rc = _lib.load().b2_u8_to_f32_norm(_lib.ptr(x), _lib.ptr(out), x.numel(), _lib.stream_ptr())
    _lib.check(rc, "b2_u8_to_f32_norm")
    _count(1)
    return out


def resize_bilinear_into(x, out, c0, factor):
    """K.resize_images(x, factor, factor, 'channels_last', 'bilinear') (model.py:357; TF 1.x legacy bilinear) of the fp32
    NHWC image x, written into channels [c0, c0+C) of the 16-bit NHWC tensor `out` (B, H*factor, W*factor, ld)."""
    B, H, W, C = x.shape
    if x.dtype != torch.float32 or out.dtype not in (torch.float16, torch.bfloat16):
        raise ValueError("resize_bilinear_into: fp32 input and a 16-bit output tensor expected")
    if tuple(out.shape[:3]) != (B, H * factor, W * factor):
        raise ValueError("resize_bilinear_into: output %s does not match %s x%d" % (tuple(out.shape), tuple(x.shape), factor))
    _require_cuda(x, out)
    rc = _lib.load().b2_resize_bilinear_into(_lib.ptr(x), B, H, W, C, factor, _lib.ptr(out), out.shape[3], c0,
                                             1 if out.dtype == torch.float16 else 0, _lib.stream_ptr())
    _lib.check(rc, "b2_resize_bilinear_into")
    _count(1)
    return out


def gate_mul(att, m, out=None, att_c0=0):
    """Multiply([sigmoid(attention), stream]) (model.py:34-36, 88-91): out = sigmoid(att[..., att_c0:att_c0+C]) * m; `att` may
    be wider than m (the attention convs of all residual blocks evaluated as one convolution)."""
    C = m.shape[-1]
    if att.dtype != m.dtype or m.dtype not in (torch.float16, torch.bfloat16) or tuple(att.shape[:-1]) != tuple(m.shape[:-1]):
        raise ValueError("gate_mul: 16-bit operands over the same pixel grid expected")
    if out is None:
        out = torch.empty_like(m)
    _require_cuda(att, m, out)
    rc = _lib.load().b2_gate_mul(_lib.ptr(att), att.shape[-1], att_c0, _lib.ptr(m), _lib.ptr(out), m.numel() // C, C,
                                 1 if m.dtype == torch.float16 else 0, _lib.stream_ptr())
    _lib.check(rc, "b2_gate_mul")
    _count(1)
    return out


def resize_nearest_bilinear_pair(x, factor):
    """Concatenate([nearest, bilinear] x`factor` resize of the fp32 NHWC image) (model.py:80-82) -> fp32 (B, fH, fW, 2C)."""
    B, H, W, C = x.shape
    if x.dtype != torch.float32:
        raise ValueError("resize_nearest_bilinear_pair expects float32")
    out = torch.empty((B, H * factor, W * factor, 2 * C), dtype=torch.float32, device=x.device)
    _require_cuda(x, out)
    rc = _lib.load().b2_resize_nearest_bilinear_pair(_lib.ptr(x), B, H, W, C, factor, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "b2_resize_nearest_bilinear_pair")
    _count(1)
    return out


def atanh_conv2d_transpose_small(x, w, bias, stride, out=None, dtype=torch.float16):
    """Conv2DTranspose(Cout, stride + 1, strides=stride, 'same')(atanh(0.99999 x)) on the 3-channel fp32 input (model.py:94-95);
    w fp32 in the Keras layout (k, k, Cout, 3)."""
    B, H, W, Cin = x.shape
    k, cout = stride + 1, w.shape[2]
    if x.dtype != torch.float32 or w.dtype != torch.float32 or tuple(w.shape) != (k, k, cout, Cin):
        raise ValueError("atanh_conv2d_transpose_small: fp32 input and a (%d,%d,Cout,%d) fp32 kernel expected" % (k, k, Cin))
    if out is None:
        out = torch.empty((B, H * stride, W * stride, cout), dtype=dtype, device=x.device)
    _require_cuda(x, w, bias, out)
    rc = _lib.load().b2_atanh_conv2d_transpose_small(_lib.ptr(x), B, H, W, Cin, _lib.ptr(w), _lib.ptr(bias), stride, cout,
                                                     _lib.ptr(out), 1 if out.dtype == torch.float16 else 0, _lib.stream_ptr())
    _lib.check(rc, "b2_atanh_conv2d_transpose_small")
    _count(1)
    return out


def f32_to_u8_denorm(x, out=None):
    """float32 [-1, 1] -> uint8: uint8(around((x + 1) * 127.5)) (data.py:253-256), bit-exact with NumPy on float32."""
    if x.dtype != torch.float32:
        raise ValueError("f32_to_u8_denorm expects float32")
    if out is None:
        out = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    _require_cuda(x, out)
    rc = _lib.load().b2_f32_to_u8_denorm(_lib.ptr(x), _lib.ptr(out), x.numel(), _lib.stream_ptr())
    _lib.check(rc, "b2_f32_to_u8_denorm")
    _count(1)
    return out


def count_saturated(x, counter):
    """counter (1-element int32 CUDA tensor) += number of elements of the 16-bit tensor x at the edge of their type
    (fp16: |x| >= 65504, inf, NaN; bf16: inf, NaN)."""
    if x.dtype not in (torch.float16, torch.bfloat16) or counter.dtype != torch.int32:
        raise ValueError("count_saturated expects a 16-bit tensor and an int32 counter")
    _require_cuda(x, counter)
    rc = _lib.load().b2_count_saturated(_lib.ptr(x), x.numel(), 1 if x.dtype == torch.float16 else 0, _lib.ptr(counter),
                                        _lib.stream_ptr())
    _lib.check(rc, "b2_count_saturated")
    _count(1)


def gather_cast(src, idx, out=None, dtype=torch.bfloat16):
    """out[i] = idx[i] >= 0 ? cast(src.flat[idx[i]]) : 0 — refreshes a packed 16-bit weight copy from fp32 masters."""
    if src.dtype != torch.float32 or idx.dtype != torch.int32:
        raise ValueError("gather_cast expects float32 source and int32 index map")
    if out is None:
        out = torch.empty(idx.shape, dtype=dtype, device=src.device)
    if out.dtype not in (torch.bfloat16, torch.float16) or out.numel() != idx.numel():
        raise ValueError("gather_cast: bad output")
    _require_cuda(src, idx, out)
    rc = _lib.load().b2_gather_cast(_lib.ptr(src), _lib.ptr(idx), idx.numel(), _lib.ptr(out),
                                    1 if out.dtype == torch.float16 else 0, _lib.stream_ptr())
    _lib.check(rc, "b2_gather_cast")
    _count(1)
    return out


# ------------------------------------------------------------------------------------------------ training step
def _dt(t):
    return {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}[t.dtype]


def _rows_c(t):
    return t.numel() // t.shape[-1], t.shape[-1]


def conv2d_s2_dgrad(dz, w_packed_t, cin, kh, kw, in_hw, out=None, gate=None, gate_slope=0.0, residual=None):
    """Data gradient of Conv2D(strides=2, 'same') (discriminator blocks, model.py:843-871): dz (B,OH,OW,Cout) ->
    (B,H,W,Cin); `w_packed_t` from conv_s2_dgrad_pack_index."""
    B, OH, OW, Cout = dz.shape
    H, W = in_hw
    f16 = _elem_flag(dz, w_packed_t)
    if out is None:
        out = torch.empty((B, H, W, cin), dtype=dz.dtype, device=dz.device)
    _require_cuda(dz, w_packed_t, out, gate, residual)
    ep = _lib.make_epilogue(out, None, None, None, ACT_NONE, 0.0, residual, f16, gate, gate_slope)
    rc = _lib.load().b2_conv2d_s2_dgrad(_lib.ptr(dz), B, OH, OW, Cout, _lib.ptr(w_packed_t), cin, kh, kw, H, W,
                                        ctypes.byref(ep), _lib.stream_ptr())
    _lib.check(rc, "b2_conv2d_s2_dgrad")
    _count(4)
    return out


def conv2d_pad(x, w_packed, cout, kh, kw, stride, pad, out_hw, out=None, out_dtype=None, shift=None, gate=None,
               gate_slope=0.0, scale=None):
    """Conv2D with explicit zero padding `pad` = (pad_h, pad_w) on an explicit output grid: padding='valid' (pad 0,
    out = (in - k) // stride + 1; make_discriminator_sparse_512, model.py:967-989) or the 'full' correlation (pad k-1)."""
    B, H, W, Cin = x.shape
    OH, OW = out_hw
    f16 = _elem_flag(x, w_packed)
    if Cin % 64 != 0 or w_packed.numel() != cout * kh * kw * Cin:
        raise ValueError("conv2d_pad: Cin must be a multiple of 64 and the packed weight [Cout][kh*kw*Cin]")
    if out is None:
        out = torch.empty((B, OH, OW, cout), dtype=out_dtype or x.dtype, device=x.device)
    _require_cuda(x, w_packed, out, shift, gate, scale)
    ep = _lib.make_epilogue(out, scale, shift, None, ACT_NONE, 0.0, None, f16, gate, gate_slope)
    rc = _lib.load().b2_conv2d_fprop_pad(_lib.ptr(x), B, H, W, Cin, _lib.ptr(w_packed), cout, kh, kw, stride, pad[0], pad[1],
                                         OH, OW, ctypes.byref(ep), _lib.stream_ptr())
    _lib.check(rc, "b2_conv2d_fprop_pad")
    _count(1)
    return out


def conv2d_strided_dgrad(dz, w_packed_t, cin, kh, kw, stride, pad, in_hw, out=None):
    """Data gradient of conv2d_pad for stride 2 or 3: dz (B,OH,OW,Cout) -> (B,H,W,Cin); `w_packed_t` from
    conv_s2_dgrad_pack_index."""
    B, OH, OW, Cout = dz.shape
    H, W = in_hw
    f16 = _elem_flag(dz, w_packed_t)
    if out is None:
        out = torch.empty((B, H, W, cin), dtype=dz.dtype, device=dz.device)
    _require_cuda(dz, w_packed_t, out)
    ep = _lib.make_epilogue(out, None, None, None, ACT_NONE, 0.0, None, f16, None, 0.0)
    rc = _lib.load().b2_conv2d_strided_dgrad(_lib.ptr(dz), B, OH, OW, Cout, _lib.ptr(w_packed_t), cin, kh, kw, stride,
                                             pad[0], pad[1], H, W, ctypes.byref(ep), _lib.stream_ptr())
    _lib.check(rc, "b2_conv2d_strided_dgrad")
    _count(stride * stride)
    return out


def conv2d_wgrad_pad(x, dz, kh, kw, stride, pad, dw):
    """conv2d_wgrad for conv2d_pad's geometry: dw (kh,kw,Cx,Cz) fp32 += sum_p x[p*stride + tap - pad] (x) dz[p]."""
    B, H, W, Cx = x.shape
    _, OH, OW, Cz = dz.shape
    if x.dtype != dz.dtype or x.dtype not in (torch.bfloat16, torch.float16):
        raise ValueError("conv2d_wgrad_pad: x and dz must share a 16-bit dtype")
    if dw.dtype != torch.float32 or dw.numel() != kh * kw * Cx * Cz:
        raise ValueError("conv2d_wgrad_pad: dw must be float32 with %d elements" % (kh * kw * Cx * Cz))
    _require_cuda(x, dz, dw)
    rc = _lib.load().b2_conv2d_wgrad_pad(_lib.ptr(x), B, H, W, Cx, _lib.ptr(dz), OH, OW, Cz, kh, kw, stride, pad[0], pad[1],
                                         _lib.ptr(dw), 1 if x.dtype == torch.float16 else 0, _lib.stream_ptr())
    _lib.check(rc, "b2_conv2d_wgrad_pad")
    _count(1)
    return dw


def conv2d_wgrad(x, dz, kh, kw, stride, dw):
    """dw (kh,kw,Cx,Cz) fp32 += sum_p x[p*stride + tap - pad] (x) dz[p]  (Conv2D: x = input, dz = output gradient;
    Conv2DTranspose: x = output gradient, dz = input, stride 2)."""
    B, H, W, Cx = x.shape
    Cz = dz.shape[-1]
    if x.dtype != dz.dtype or x.dtype not in (torch.bfloat16, torch.float16):
        raise ValueError("conv2d_wgrad: x and dz must share a 16-bit dtype")
    if tuple(dz.shape[:3]) != (B, tf_same_out(H, stride), tf_same_out(W, stride)):
        raise ValueError("conv2d_wgrad: dz %s is not the stride-%d SAME output grid of x %s" % (tuple(dz.shape), stride, tuple(x.shape)))
    if dw.dtype != torch.float32 or dw.numel() != kh * kw * Cx * Cz:
        raise ValueError("conv2d_wgrad: dw must be float32 with %d elements" % (kh * kw * Cx * Cz))
    _require_cuda(x, dz, dw)
    rc = _lib.load().b2_conv2d_wgrad(_lib.ptr(x), B, H, W, Cx, _lib.ptr(dz), Cz, kh, kw, stride, _lib.ptr(dw),
                                     1 if x.dtype == torch.float16 else 0, _lib.stream_ptr())
    _lib.check(rc, "b2_conv2d_wgrad")
    _count(1)
    return dw


def colstats(x, ssum, ssq=None):
    rows, C = _rows_c(x)
    _require_cuda(x, ssum, ssq)
    rc = _lib.load().b2_colstats(_lib.ptr(x), rows, C, _dt(x), _lib.ptr(ssum), _lib.ptr(ssq), _lib.stream_ptr())
    _lib.check(rc, "b2_colstats")
    _count(1)


def bn_finalize(ssum, ssq, count, gamma, beta, scale, shift, mean=None, invstd=None, moving_mean=None, moving_var=None,
                eps=1e-3, momentum=0.99):
    rc = _lib.load().b2_bn_finalize(_lib.ptr(ssum), _lib.ptr(ssq), float(count), _lib.ptr(gamma), _lib.ptr(beta), eps, momentum,
                                    _lib.ptr(moving_mean), _lib.ptr(moving_var), _lib.ptr(scale), _lib.ptr(shift),
                                    _lib.ptr(mean), _lib.ptr(invstd), gamma.numel(), _lib.stream_ptr())
    _lib.check(rc, "b2_bn_finalize")
    _count(1)


def bn_fold(gamma, beta, mean, var, bias, scale, shift, eps=1e-3):
    rc = _lib.load().b2_bn_fold(_lib.ptr(gamma), _lib.ptr(beta), _lib.ptr(mean), _lib.ptr(var), _lib.ptr(bias), eps,
                                _lib.ptr(scale), _lib.ptr(shift), gamma.numel(), _lib.stream_ptr())
    _lib.check(rc, "b2_bn_fold")
    _count(1)


def affine_act(z, out, scale=None, shift=None, alpha=None, act=ACT_NONE, leaky=0.0, residual=None):
    rows, C = _rows_c(z)
    _require_cuda(z, out, residual)
    rc = _lib.load().b2_affine_act(_lib.ptr(z), rows, C, _dt(z), _lib.ptr(scale), _lib.ptr(shift), _lib.ptr(alpha), act, leaky,
                                   _lib.ptr(residual), _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "b2_affine_act")
    _count(1)
    return out


def bn_bwd_reduce(z, g, scale, shift, alpha, act, leaky, mean, invstd, s1, s2, dalpha=None):
    rows, C = _rows_c(z)
    _require_cuda(z, g)
    rc = _lib.load().b2_bn_bwd_reduce(_lib.ptr(z), _lib.ptr(g), rows, C, _dt(z), _lib.ptr(scale), _lib.ptr(shift),
                                      _lib.ptr(alpha), act, leaky, _lib.ptr(mean), _lib.ptr(invstd), _lib.ptr(s1), _lib.ptr(s2),
                                      _lib.ptr(dalpha), _lib.stream_ptr())
    _lib.check(rc, "b2_bn_bwd_reduce")
    _count(1)


def bn_bwd_apply(z, g, scale, shift, alpha, act, leaky, mean, invstd, s1, s2, count, dz):
    rows, C = _rows_c(z)
    _require_cuda(z, g, dz)
    rc = _lib.load().b2_bn_bwd_apply(_lib.ptr(z), _lib.ptr(g), rows, C, _dt(z), _lib.ptr(scale), _lib.ptr(shift), _lib.ptr(alpha),
                                     act, leaky, _lib.ptr(mean), _lib.ptr(invstd), _lib.ptr(s1), _lib.ptr(s2), float(count),
                                     _lib.ptr(dz), _lib.stream_ptr())
    _lib.check(rc, "b2_bn_bwd_apply")
    _count(1)
    return dz


_bn_rows = None


def bn_part_rows():
    """Rows a stage-1 BatchNorm statistics producer may write (2 per SM)."""
    global _bn_rows
    if _bn_rows is None:
        _bn_rows = int(_lib.load().b2_bn_part_rows())
    return _bn_rows


def sm_count():
    return int(_lib.load().b2_sm_count())


def _peer_args(peer):
    """(regions, rank, world, epoch) of one exchange on `peer` (upscaler.peer.PeerReducer), or the inactive form."""
    if peer is None:
        return None, 0, 1, 0, None
    epoch = peer.next_epoch()
    return peer.regions, peer.rank, peer.world, epoch, peer.epoch_base_ptr()


def bn_stats_partial(x, part=None):
    """Stage 1 of the train-mode BatchNorm statistics: per-CTA partial [sum | sum of squares] rows of x viewed as
    [rows][C].  Returns (part, nparts)."""
    rows, C = _rows_c(x)
    if part is None:
        part = torch.empty((bn_part_rows(), 2 * C), dtype=torch.float32, device=x.device)
    _require_cuda(x, part)
    n = ctypes.c_int(0)
    rc = _lib.load().b2_bn_stats_partial(_lib.ptr(x), rows, C, _dt(x), _lib.ptr(part), ctypes.byref(n), _lib.stream_ptr())
    _lib.check(rc, "b2_bn_stats_partial")
    _count(1)
    return part, n.value


def bn_fwd_finalize(part, nparts, C, count, gamma, beta, aff, moving_mean=None, moving_var=None, peer=None, sums_out=None,
                    eps=1e-3, momentum=0.99, range_watch=None, range_limit=0.0):
    """Stage 2: partial rows (+ the other replicas' through `peer`) -> aff = [scale | shift | mean | invstd]; updates the
    moving statistics when given.  aff=None with sums_out: only the summed statistics (NCCL fallback).  `range_watch`
    (one float32) is set to 1 when a channel's root mean square reaches `range_limit` (guard of fp16 storage)."""
    regions, rank, world, epoch, ebase = _peer_args(peer)
    rc = _lib.load().b2_bn_fwd_finalize(_lib.ptr(part), int(nparts), int(C), float(count), regions, rank, world, epoch, ebase,
                                        _lib.ptr(sums_out), _lib.ptr(gamma), _lib.ptr(beta), eps, momentum,
                                        _lib.ptr(moving_mean), _lib.ptr(moving_var), _lib.ptr(aff), _lib.ptr(range_watch),
                                        float(range_limit), _lib.stream_ptr())
    _lib.check(rc, "b2_bn_fwd_finalize")
    _count(1)
    return aff


def bn_fwd_stats_static(x, count, gamma=None, beta=None, sums_out=None, aff=None, eps=1e-3):
    """Forward BatchNorm statistics without a scratch tensor: b2_bn_fwd_stats_static keeps the partial rows in an array
    built into the library (C <= 512), for embedding hosts that may not allocate device memory.  This wrapper allocates
    nothing either: the caller passes the outputs (sums_out [2*C] and / or aff [4*C]).  Calls on one device must not overlap."""
    rows, C = _rows_c(x)
    _require_cuda(x)
    rc = _lib.load().b2_bn_fwd_stats_static(_lib.ptr(x), rows, C, _dt(x), float(count), _lib.ptr(gamma), _lib.ptr(beta), float(eps),
                                            _lib.ptr(sums_out), _lib.ptr(aff), _lib.stream_ptr())
    _lib.check(rc, "b2_bn_fwd_stats_static")
    _count(2)
    return sums_out, aff


def bn_bwd_partial(z, g, aff, alpha, act, leaky, part=None):
    """Stage 1 of the BatchNorm backward sums through act(BN(z)): partial [sum dy | sum dy*xhat | sum g*min(u,0)] rows."""
    rows, C = _rows_c(z)
    if part is None:
        part = torch.empty((bn_part_rows(), 3 * C), dtype=torch.float32, device=z.device)
    _require_cuda(z, g, aff, part)
    n = ctypes.c_int(0)
    rc = _lib.load().b2_bn_bwd_partial(_lib.ptr(z), _lib.ptr(g), rows, C, _dt(z), _lib.ptr(aff), _lib.ptr(alpha), act, leaky,
                                       _lib.ptr(part), ctypes.byref(n), _lib.stream_ptr())
    _lib.check(rc, "b2_bn_bwd_partial")
    _count(1)
    return part, n.value


def bn_bwd_finalize(part, nparts, C, sums, dgamma=None, dbeta=None, dalpha=None, gscale=1.0, peer=None, raw_out=None):
    """Stage 2: sums = [s1 | s2] over all replicas; dbeta += s1*gscale, dgamma += s2*gscale, dalpha += local third sum."""
    regions, rank, world, epoch, ebase = _peer_args(peer)
    rc = _lib.load().b2_bn_bwd_finalize(_lib.ptr(part), int(nparts), int(C), regions, rank, world, epoch, ebase, _lib.ptr(raw_out),
                                        _lib.ptr(sums), _lib.ptr(dgamma), _lib.ptr(dbeta), _lib.ptr(dalpha), float(gscale),
                                        _lib.stream_ptr())
    _lib.check(rc, "b2_bn_bwd_finalize")
    _count(1)
    return sums


def maxpool2x2(x, out=None):
    B, H, W, C = x.shape
    if out is None:
        out = torch.empty((B, H // 2, W // 2, C), dtype=x.dtype, device=x.device)
    _require_cuda(x, out)
    rc = _lib.load().b2_maxpool2x2(_lib.ptr(x), B, H, W, C, _dt(x), _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "b2_maxpool2x2")
    _count(1)
    return out


def maxpool2x2_bwd(x, y, dy, relu_gate=False, out=None):
    B, H, W, C = x.shape
    if out is None:
        out = torch.empty_like(x)
    _require_cuda(x, y, dy, out)
    rc = _lib.load().b2_maxpool2x2_bwd(_lib.ptr(x), _lib.ptr(y), _lib.ptr(dy), B, H, W, C, _dt(x), 1 if relu_gate else 0,
                                       _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "b2_maxpool2x2_bwd")
    _count(1)
    return out


def loss_pair(a, b, loss, mae=False, weight=1.0, gscale=1.0, grad=None):
    """loss (1-element fp32 tensor) += weight * mean(f(a-b)); grad = gscale * f'(a-b)/n (same dtype as a)."""
    if a.dtype != b.dtype or a.numel() != b.numel():
        raise ValueError("loss_pair: operands must match")
    _require_cuda(a, b, loss, grad)
    rc = _lib.load().b2_loss_pair(_lib.ptr(a), _lib.ptr(b), a.numel(), _dt(a), 1 if mae else 0, weight, gscale, _lib.ptr(loss),
                                  _lib.ptr(grad), _lib.stream_ptr())
    _lib.check(rc, "b2_loss_pair")
    _count(1)


SCORE_ACTS = {"sigmoid": 1, "log-sigm": 2, "tanh": 3, "bi-log": 4}


def score_act(z, kind, dy=None, out=None):
    """Discriminator output activation (model.py:885-892) of the fp32 scores z; with dy: its backward dy * f'(z)."""
    if out is None:
        out = torch.empty_like(z)
    _require_cuda(z, out, dy)
    rc = _lib.load().b2_score_act(_lib.ptr(z), _lib.ptr(dy), _lib.ptr(out), z.numel(), SCORE_ACTS[kind], _lib.stream_ptr())
    _lib.check(rc, "b2_score_act")
    _count(1)
    return out


def mean_prod(x, y, loss, weight=1.0, gscale=1.0, grad=None, yconst=1.0):
    rc = _lib.load().b2_mean_prod(_lib.ptr(x), _lib.ptr(y), yconst, x.numel(), weight, gscale, _lib.ptr(loss), _lib.ptr(grad),
                                  _lib.stream_ptr())
    _lib.check(rc, "b2_mean_prod")
    _count(1)


def tanh_bwd(y, g1, dz, g2=None, g3=None):
    rc = _lib.load().b2_tanh_bwd(_lib.ptr(y), _lib.ptr(g1), _lib.ptr(g2), _lib.ptr(g3), y.numel(), _lib.ptr(dz), _lib.stream_ptr())
    _lib.check(rc, "b2_tanh_bwd")
    _count(1)
    return dz


def gather_f32(src, idx, out, accumulate=False):
    rc = _lib.load().b2_gather_f32(_lib.ptr(src), _lib.ptr(idx), idx.numel(), _lib.ptr(out), 1 if accumulate else 0,
                                   _lib.stream_ptr())
    _lib.check(rc, "b2_gather_f32")
    _count(1)
    return out


def colsum_f32(x, out):
    """out[c] += sum over rows of the fp32 tensor x (..., C), C <= 8."""
    rows, C = _rows_c(x)
    rc = _lib.load().b2_colsum_f32(_lib.ptr(x), rows, C, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "b2_colsum_f32")
    _count(1)


def dense_fwd(x, W, b, out):
    n, K = x.shape[0], x.numel() // x.shape[0]
    N = W.shape[1]
    rc = _lib.load().b2_dense_fwd(_lib.ptr(x), _dt(x), n, K, N, _lib.ptr(W), _lib.ptr(b), _lib.ptr(out), _dt(out), _lib.stream_ptr())
    _lib.check(rc, "b2_dense_fwd")
    _count(1)
    return out


def dense_bwd_data(dy, W, dx):
    n, N = dy.shape[0], W.shape[1]
    K = W.shape[0]
    rc = _lib.load().b2_dense_bwd_data(_lib.ptr(dy), _dt(dy), n, K, N, _lib.ptr(W), _lib.ptr(dx), _dt(dx), _lib.stream_ptr())
    _lib.check(rc, "b2_dense_bwd_data")
    _count(1)
    return dx


def dense_bwd_weight(x, dy, dW, db):
    n, K, N = x.shape[0], dW.shape[0], dW.shape[1]
    rc = _lib.load().b2_dense_bwd_weight(_lib.ptr(x), _dt(x), _lib.ptr(dy), _dt(dy), n, K, N, _lib.ptr(dW), _lib.ptr(db),
                                         _lib.stream_ptr())
    _lib.check(rc, "b2_dense_bwd_weight")
    _count(1)


def cast_scale(x, out, scale=1.0):
    rc = _lib.load().b2_cast_scale(_lib.ptr(x), _dt(x), x.numel(), scale, _lib.ptr(out), _dt(out), _lib.stream_ptr())
    _lib.check(rc, "b2_cast_scale")
    _count(1)
    return out


def adam(params, grads, m, v, lr_t, beta1, beta2, eps, gscale=1.0, guards=()):
    """Keras Adam over a flat fp32 buffer.  `guards`: up to two int32 CUDA tensors of equal length; the update is dropped on
    the device when any of their words is non-zero (b2_adam_guarded)."""
    guards = [g for g in guards if g is not None]
    if not guards:
        rc = _lib.load().b2_adam(_lib.ptr(params), _lib.ptr(grads), _lib.ptr(m), _lib.ptr(v), params.numel(), lr_t, beta1, beta2, eps,
                                 gscale, _lib.stream_ptr())
        _lib.check(rc, "b2_adam")
    else:
        if len(guards) > 2 or any(g.dtype != torch.int32 or g.numel() != guards[0].numel() for g in guards):
            raise ValueError("adam: at most two int32 guard tensors of equal length")
        _require_cuda(*guards)
        g1 = guards[1] if len(guards) > 1 else None
        rc = _lib.load().b2_adam_guarded(_lib.ptr(params), _lib.ptr(grads), _lib.ptr(m), _lib.ptr(v), params.numel(), lr_t, beta1, beta2,
                                         eps, gscale, _lib.ptr(guards[0]), _lib.ptr(g1), guards[0].numel(), _lib.stream_ptr())
        _lib.check(rc, "b2_adam_guarded")
    _count(1)


# ------------------------------------------------------------------------------------------------ packing index maps
FINAL_TAPS_PER_GROUP = 21      # 21 taps x 3 channels = 63 of the 64 channels of one fold group


def _final_fold_channel(t, co, cout):
    return 64 * (t // FINAL_TAPS_PER_GROUP) + (t % FINAL_TAPS_PER_GROUP) * cout + co


def final_dgrad_pack_index(kh, kw, cin, cout, chan_map=None):
    """Data gradient of the final Conv2D(3, 9x9) (model.py:290) as a 1x1 conv over the tap-folded output gradient F
    (F[q, f(t,co)] = dz[q - d(t), co]): packed [rows = our input channels][256 folded channels], entry = W[t, c, co].
    `chan_map`: our channel j is Keras input channel chan_map[j] (-1 = padding), see conv_scatter_pack_index."""
    T = kh * kw
    cm = np.arange(cin) if chan_map is None else np.asarray(chan_map, np.int64)
    idx = -np.ones((cm.size, 64 * (-(-T // FINAL_TAPS_PER_GROUP))), np.int64)
    for t in range(T):
        for co in range(cout):
            idx[:, _final_fold_channel(t, co, cout)] = np.where(cm >= 0, (t * cin + cm) * cout + co, -1)
    return idx.astype(np.int32)


def final_wgrad_gather_index(kh, kw, cin, cout, chan_map=None):
    """Keras (kh,kw,Cin,Cout) gradient of the final conv <- raw[j][f] = sum_p X[p,j] * F[p,f] (raw row pitch 256), j = our
    channel holding Keras channel c (the inverse of `chan_map`)."""
    T = kh * kw
    nf = 64 * (-(-T // FINAL_TAPS_PER_GROUP))
    idx = np.zeros((T, cin, cout), np.int64)
    if chan_map is None:
        row = np.arange(cin)
    else:
        cm = np.asarray(chan_map, np.int64)
        row = np.zeros(cin, np.int64)
        row[cm[cm >= 0]] = np.nonzero(cm >= 0)[0]
    c = row[:, None]
    for t in range(T):
        f = np.array([_final_fold_channel(t, co, cout) for co in range(cout)])[None, :]
        idx[t] = c * nf + f
    return idx.astype(np.int32)


def padded_gather_index(true_shape, padded_shape):
    """Flat positions, in a tensor of `padded_shape`, of the leading `true_shape` block (same rank): the weight gradient of
    a layer run with zero-padded channels is produced in the padded layout and gathered back into the Keras one."""
    grids = np.meshgrid(*[np.arange(n) for n in true_shape], indexing="ij")
    flat = np.zeros(true_shape, np.int64)
    for g, n in zip(grids, padded_shape):
        flat = flat * n + g
    return flat.astype(np.int32)


def rowfold_wgrad_gather_index(kh, kw, cin, cout):
    """Keras (kh,kw,Cin,Cout) gradient of a row-folded narrow conv <- raw (kh,1,64,Cout) from conv2d_wgrad(kh x 1) on
    the folded input (channel = kx*Cin + c)."""
    ky = np.arange(kh)[:, None, None, None]
    kx = np.arange(kw)[None, :, None, None]
    c = np.arange(cin)[None, None, :, None]
    co = np.arange(cout)[None, None, None, :]
    return ((ky * 64 + kx * cin + c) * cout + co).astype(np.int32)


def fullfold_wgrad_gather_index(kh, kw, cin, cout):
    """Keras (kh,kw,Cin,Cout) gradient of a fully folded narrow conv <- raw (1,1,64,Cout): the first kh*kw*Cin rows."""
    return np.arange(kh * kw * cin * cout, dtype=np.int32).reshape(kh, kw, cin, cout)


def conv_dgrad_pack_index(kh, kw, cin, cout):
    """Keras Conv2D kernel (kh,kw,Cin,Cout) -> packed [Cin][kh*kw (flipped)][Cout]: the stride-1 data gradient
    dX = conv(dZ, flip(W)^T) runs through conv2d with this weight."""
    T = kh * kw
    t = np.arange(T)
    tf = (kh - 1 - t // kw) * kw + (kw - 1 - t % kw)
    ci = np.arange(cin)[:, None, None]
    co = np.arange(cout)[None, None, :]
    return ((tf[None, :, None] * cin + ci) * cout + co).astype(np.int32)


def conv_s2_dgrad_pack_index(kh, kw, cin, cout):
    """Keras Conv2D kernel (kh,kw,Cin,Cout) -> packed [Cin][kh*kw][Cout] (taps NOT flipped) for conv2d_s2_dgrad."""
    t = np.arange(kh * kw)[None, :, None]
    ci = np.arange(cin)[:, None, None]
    co = np.arange(cout)[None, None, :]
    return ((t * cin + ci) * cout + co).astype(np.int32)


def convT_dgrad_pack_index(kh, kw, cin, cout, cin_pad=None, cout_pad=None):
    """Keras Conv2DTranspose kernel (kh,kw,Cout,Cin) -> packed [cin_pad][kh*kw][cout_pad] (-1 = zero): the data gradient
    of Conv2DTranspose(stride 2) is a stride-2 'same' conv2d of the output gradient with this weight."""
    t = np.arange(kh * kw)[None, :, None]
    ci = np.arange(cin)[:, None, None]
    co = np.arange(cout)[None, None, :]
    idx = ((t * cout + co) * cin + ci).astype(np.int64)
    if (cin_pad or cin) == cin and (cout_pad or cout) == cout:
        return idx.astype(np.int32)
    full = -np.ones((cin_pad or cin, kh * kw, cout_pad or cout), np.int64)
    full[:cin, :, :cout] = idx
    return full.astype(np.int32)



def conv_pack_index(kh, kw, cin, cout, cin_pad=None):
    """Keras Conv2D kernel (kh,kw,Cin,Cout) -> packed [Cout][kh*kw][cin_pad] index map (int32, -1 = zero)."""
    cp = cin_pad or cin
    T = kh * kw
    idx = -np.ones((cout, T, cp), np.int64)
    t = np.arange(T)[None, :, None]
    c = np.arange(cin)[None, None, :]
    co = np.arange(cout)[:, None, None]
    idx[:, :, :cin] = (t * cin + c) * cout + co
    return idx.astype(np.int32)


def convT_pack_index(kh, kw, cin, cout, cin_pad=None, cout_pad=None):
    """Keras Conv2DTranspose kernel (kh,kw,Cout,Cin) -> packed [cout_pad][kh*kw][cin_pad] (-1 = zero: padded channels)."""
    T = kh * kw
    t = np.arange(T)[None, :, None]
    c = np.arange(cin)[None, None, :]
    co = np.arange(cout)[:, None, None]
    idx = ((t * cout + co) * cin + c).astype(np.int64)
    if (cin_pad or cin) == cin and (cout_pad or cout) == cout:
        return idx.astype(np.int32)
    full = -np.ones((cout_pad or cout, T, cin_pad or cin), np.int64)
    full[:cout, :, :cin] = idx
    return full.astype(np.int32)


def conv_rowfold_pack_index(kh, kw, cin, cout):
    """Narrow-Cin conv whose kw taps of each filter row are folded into channels (fold_taps with dx taps):
    packed [Cout][kh][64], channel = kx*Cin + c."""
    if kw * cin > 64:
        raise ValueError("row fold needs kw*Cin <= 64")
    idx = -np.ones((cout, kh, 64), np.int64)
    ky = np.arange(kh)[None, :, None, None]
    kx = np.arange(kw)[None, None, :, None]
    c = np.arange(cin)[None, None, None, :]
    co = np.arange(cout)[:, None, None, None]
    src = ((ky * kw + kx) * cin + c) * cout + co
    idx[:, :, :kw * cin] = src.reshape(cout, kh, kw * cin)
    return idx.astype(np.int32)


def conv_fullfold_pack_index(kh, kw, cin, cout):
    """Narrow-Cin conv with ALL taps folded into channels: packed [Cout][1][64], channel = (ky*kw+kx)*Cin + c."""
    if kh * kw * cin > 64:
        raise ValueError("full fold needs kh*kw*Cin <= 64")
    idx = -np.ones((cout, 1, 64), np.int64)
    t = np.arange(kh * kw)[None, :, None]
    c = np.arange(cin)[None, None, :]
    co = np.arange(cout)[:, None, None]
    idx[:, 0, :kh * kw * cin] = ((t * cin + c) * cout + co).reshape(cout, kh * kw * cin)
    return idx.astype(np.int32)


def conv_scatter_pack_index(kh, kw, cin, cout, chan_map=None):
    """Keras Conv2D kernel (kh,kw,Cin,Cout) -> packed [256][Cin'] for conv2d_scatter: row n = (dy*kw + dx)*Cout + co.
    `chan_map` (length Cin', optional): our input channel j reads Keras input channel chan_map[j] (-1 = zero padding) --
    used when the tensor the conv reads is a padded / permuted form of the Keras one (skip-con's Concatenate)."""
    if kh * kw * cout > 256:
        raise ValueError("conv_scatter needs kh*kw*Cout <= 256")
    cm = np.arange(cin) if chan_map is None else np.asarray(chan_map, np.int64)
    idx = -np.ones((256, cm.size), np.int64)
    t = np.arange(kh * kw)[:, None, None]
    co = np.arange(cout)[None, :, None]
    c = cm[None, None, :]
    full = np.where(c >= 0, (t * cin + c) * cout + co, -1)
    idx[:kh * kw * cout] = full.reshape(kh * kw * cout, cm.size)
    return idx.astype(np.int32)


def conv_rows_pack_index(kh, kw, cin, cout):
    """Keras Conv2D kernel (kh,kw,Cin,Cout) -> packed [32][kw*Cin] for conv2d_rows: row n = dy*Cout + co,
    column k = dx*Cin + c."""
    if kh * cout > 32:
        raise ValueError("conv_rows needs kh*Cout <= 32")
    idx = -np.ones((32, kw, cin), np.int64)
    dy = np.arange(kh)[:, None, None, None]
    co = np.arange(cout)[None, :, None, None]
    dx = np.arange(kw)[None, None, :, None]
    c = np.arange(cin)[None, None, None, :]
    src = ((dy * kw + dx) * cin + c) * cout + co
    idx[:kh * cout] = src.reshape(kh * cout, kw, cin)
    return idx.astype(np.int32)
