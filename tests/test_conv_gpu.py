"""-m gpu parity of the individual convolution entry points (through the C ABI) against the CPU oracle's
TensorFlow-semantics convolutions, on the same 16-bit-rounded operands (so the only difference is fp32 accumulation
order): tolerance 2e-3 absolute + 1e-2 relative on bf16/fp16 outputs, 1e-3 on fp32 outputs.
"""
import numpy as np
import pytest
import torch

from oracle import srgan_oracle as O

pytestmark = pytest.mark.gpu


def _rnd(shape, rng, scale=1.0):
    return (rng.standard_normal(shape) * scale).astype(np.float32)


def _q(a, dtype):
    return torch.from_numpy(a).to(dtype).to(torch.float32)


def _close(got, ref, atol, rtol):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    err = np.abs(got - ref)
    bad = err > atol + rtol * np.abs(ref)
    assert not bad.any(), "max err %.5f at %s (ref %.4f)" % (err.max(), np.unravel_index(err.argmax(), err.shape),
                                                              ref.flat[err.argmax()])


CONV_CASES = [
    # B, H, W, Cin, Cout, k, stride
    (1, 20, 33, 64, 64, 3, 1),
    (2, 17, 40, 64, 64, 5, 1),
    (1, 7, 5, 64, 64, 3, 1),       # image smaller than one tile
    (2, 24, 24, 64, 128, 3, 1),
    (1, 12, 12, 256, 512, 3, 1),
    (2, 32, 32, 64, 128, 3, 2),    # even input, stride 2: TF pads (0,1)
    (1, 27, 15, 128, 256, 3, 2),   # odd input, stride 2: TF pads (1,1)
    (3, 2, 2, 512, 512, 3, 2),     # discriminator tail: 2x2 -> 1x1
]


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("case", CONV_CASES)
def test_conv2d_fprop(case, dtype):
    from upscaler import ops
    B, H, W, Cin, Cout, k, s = case
    rng = np.random.default_rng(hash(case) % 2 ** 31)
    x = _rnd((B, H, W, Cin), rng)
    kern = _rnd((k, k, Cin, Cout), rng, 1.0 / np.sqrt(k * k * Cin))
    scale = rng.uniform(0.5, 1.5, Cout).astype(np.float32)
    shift = _rnd((Cout,), rng, 0.1)
    alpha = rng.uniform(0, 0.3, Cout).astype(np.float32)
    xq, kq = _q(x, dtype), _q(kern, dtype)
    ref = O.conv2d_same(xq.permute(0, 3, 1, 2), kq, None, stride=s)
    ref = ref * torch.from_numpy(scale).view(1, -1, 1, 1) + torch.from_numpy(shift).view(1, -1, 1, 1)
    ref = O.prelu(ref, torch.from_numpy(alpha)).permute(0, 2, 3, 1).numpy()
    idx = torch.from_numpy(ops.conv_pack_index(k, k, Cin, Cout)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern).cuda(), idx, dtype=dtype)
    y = ops.conv2d(torch.from_numpy(x).cuda().to(dtype), wp, Cout, k, k, stride=s, scale=torch.from_numpy(scale).cuda(),
                   shift=torch.from_numpy(shift).cuda(), alpha=torch.from_numpy(alpha).cuda(), act=ops.ACT_PRELU)
    torch.cuda.synchronize()
    assert tuple(y.shape) == ref.shape
    _close(y.float().cpu().numpy(), ref, 4e-3 if dtype == torch.float16 else 2e-2, 1e-2)


@pytest.mark.parametrize("k,Cin,H,W", [(3, 64, 12, 20), (5, 64, 9, 7), (3, 256, 10, 16), (5, 256, 6, 6)])
def test_conv2d_transpose_s2(k, Cin, H, W):
    from upscaler import ops
    dtype = torch.float16
    rng = np.random.default_rng(k * 1000 + Cin)
    x = _rnd((2, H, W, Cin), rng)
    kern = _rnd((k, k, 256, Cin), rng, 1.0 / np.sqrt(k * k * Cin / 4))
    bias = _rnd((256,), rng, 0.1)
    ref = O.leaky(O.conv2d_transpose_same_s2(_q(x, dtype).permute(0, 3, 1, 2), _q(kern, dtype), torch.from_numpy(bias)),
                  0.2).permute(0, 2, 3, 1).numpy()
    idx = torch.from_numpy(ops.convT_pack_index(k, k, Cin, 256)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern).cuda(), idx, dtype=dtype)
    y = ops.conv2d_transpose_s2(torch.from_numpy(x).cuda().to(dtype), wp, 256, k, k, shift=torch.from_numpy(bias).cuda(),
                                act=ops.ACT_LEAKY, leaky=0.2)
    torch.cuda.synchronize()
    assert tuple(y.shape) == (2, 2 * H, 2 * W, 256)
    _close(y.float().cpu().numpy(), ref, 4e-3, 1e-2)


@pytest.mark.parametrize("k,Cin,B,H,W", [(3, 64, 2, 32, 24), (3, 256, 1, 48, 24), (3, 256, 3, 47, 23), (5, 64, 1, 33, 31),
                                         (3, 128, 5, 16, 8)])
@pytest.mark.parametrize("pair", [True, False])
def test_conv2d_transpose_s2_fused_phases(k, Cin, B, H, W, pair):
    """Images large enough for the fused-phase kernels (b2_conv_phases2.cu over CTA pairs, b2_conv_phases.cu single CTA):
    odd tile counts (the last pair's second CTA idles), ragged right / bottom edges, TMA-store epilogue."""
    from upscaler import ops
    dtype = torch.float16
    rng = np.random.default_rng(k * 1000 + Cin + H)
    x = _rnd((B, H, W, Cin), rng)
    kern = _rnd((k, k, 256, Cin), rng, 1.0 / np.sqrt(k * k * Cin / 4))
    bias = _rnd((256,), rng, 0.1)
    ref = O.leaky(O.conv2d_transpose_same_s2(_q(x, dtype).permute(0, 3, 1, 2), _q(kern, dtype), torch.from_numpy(bias)),
                  0.2).permute(0, 2, 3, 1).numpy()
    idx = torch.from_numpy(ops.convT_pack_index(k, k, Cin, 256)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern).cuda(), idx, dtype=dtype)
    ops.set_pair_kernels(pair)
    try:
        y = ops.conv2d_transpose_s2(torch.from_numpy(x).cuda().to(dtype), wp, 256, k, k, shift=torch.from_numpy(bias).cuda(),
                                    act=ops.ACT_LEAKY, leaky=0.2)
        torch.cuda.synchronize()
    finally:
        ops.set_pair_kernels(True)
    assert tuple(y.shape) == (B, 2 * H, 2 * W, 256)
    _close(y.float().cpu().numpy(), ref, 4e-3, 1e-2)


@pytest.mark.parametrize("shape", [(1, 24, 40), (2, 9, 150), (1, 3, 5), (1, 20, 128), (1, 31, 257)])
def test_final_conv_rows(shape):
    """final/conv 9x9 256->3 + tanh, row-streaming kernel, ragged widths (strip edges) and heights < filter."""
    from upscaler import ops
    dtype = torch.float16
    B, H, W = shape
    rng = np.random.default_rng(H * 1000 + W)
    x = _rnd((B, H, W, 256), rng)
    kern = _rnd((9, 9, 256, 3), rng, 1.0 / np.sqrt(81 * 256))
    bias = _rnd((3,), rng, 0.1)
    ref = torch.tanh(O.conv2d_same(_q(x, dtype).permute(0, 3, 1, 2), _q(kern, dtype), torch.from_numpy(bias)))
    ref = ref.permute(0, 2, 3, 1).numpy()
    idx = torch.from_numpy(ops.conv_rows_pack_index(9, 9, 256, 3)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern).cuda(), idx, dtype=dtype)
    y = ops.conv2d_rows(torch.from_numpy(x).cuda().to(dtype), wp, 3, 9, 9, bias=torch.from_numpy(bias).cuda())
    torch.cuda.synchronize()
    _close(y.cpu().numpy(), ref, 1e-3, 1e-3)
    # the generic implicit-GEMM path must agree too
    idx2 = torch.from_numpy(ops.conv_pack_index(9, 9, 256, 3)).cuda()
    wp2 = ops.gather_cast(torch.from_numpy(kern).cuda(), idx2, dtype=dtype)
    y2 = ops.conv2d(torch.from_numpy(x).cuda().to(dtype), wp2, 3, 9, 9, out_dtype=torch.float32,
                    shift=torch.from_numpy(bias).cuda(), act=ops.ACT_TANH)
    torch.cuda.synchronize()
    _close(y2.cpu().numpy(), ref, 1e-3, 1e-3)


def test_head_conv_folded():
    """initial/conv 9x9 3->64: horizontal taps folded into channels, then a 9x1 conv."""
    from upscaler import ops
    dtype = torch.float16
    rng = np.random.default_rng(5)
    x = rng.uniform(-1, 1, (2, 19, 23, 3)).astype(np.float32)
    kern = _rnd((9, 9, 3, 64), rng, 1.0 / np.sqrt(243))
    bias = _rnd((64,), rng, 0.1)
    ref = O.conv2d_same(_q(x, dtype).permute(0, 3, 1, 2), _q(kern, dtype), torch.from_numpy(bias))
    ref = ref.permute(0, 2, 3, 1).numpy()
    xp = ops.fold_taps(torch.from_numpy(x).cuda(), [(0, dx - 4) for dx in range(9)], dtype=dtype)
    idx = torch.from_numpy(ops.conv_rowfold_pack_index(9, 9, 3, 64)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern).cuda(), idx, dtype=dtype)
    y = ops.conv2d(xp, wp, 64, 9, 1, shift=torch.from_numpy(bias).cuda())
    torch.cuda.synchronize()
    _close(y.float().cpu().numpy(), ref, 4e-3, 1e-2)


def test_direct_crosscheck_matches_tensor_core_path():
    from upscaler import ops
    rng = np.random.default_rng(11)
    x = torch.from_numpy(_rnd((1, 16, 24, 64), rng)).cuda().to(torch.bfloat16)
    kern = torch.from_numpy(_rnd((3, 3, 64, 64), rng, 0.05)).cuda()
    wp = ops.gather_cast(kern, torch.from_numpy(ops.conv_pack_index(3, 3, 64, 64)).cuda())
    a = ops.conv2d(x, wp, 64, 3, 3, act=ops.ACT_RELU)
    b = ops.conv2d(x, wp, 64, 3, 3, act=ops.ACT_RELU, direct=True)
    torch.cuda.synchronize()
    _close(a.float().cpu().numpy(), b.float().cpu().numpy(), 2e-2, 1e-2)


def test_errors_are_loud():
    from upscaler import ops
    x = torch.zeros((1, 8, 8, 3), dtype=torch.bfloat16, device="cuda")
    w = torch.zeros((64 * 9 * 3,), dtype=torch.bfloat16, device="cuda")
    with pytest.raises(ValueError):
        ops.conv2d(x, w, 64, 3, 3)  # Cin not a multiple of 64
    with pytest.raises(ValueError):
        ops.conv2d(torch.zeros((1, 8, 8, 64), dtype=torch.float16, device="cuda"), w, 64, 3, 3)  # dtype mismatch


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(1, 24, 40, 256), (2, 9, 150, 256), (1, 3, 5, 256), (1, 20, 96, 128), (1, 31, 257, 64),
                                   (3, 12, 193, 256)])
def test_final_conv_scatter(shape, dtype):
    """final/conv 9x9 Cin->3 + tanh, scatter-form kernel (all 81 taps in the GEMM N dimension, shuffle + register-ring
    epilogue): ragged widths around the 96-pixel strip, heights below the filter size, several images (segments of
    the flattened row space crossing image boundaries)."""
    from upscaler import ops
    B, H, W, C = shape
    rng = np.random.default_rng(H * 1000 + W)
    x = _rnd((B, H, W, C), rng)
    kern = _rnd((9, 9, C, 3), rng, 1.0 / np.sqrt(81 * C))
    bias = _rnd((3,), rng, 0.1)
    ref = torch.tanh(O.conv2d_same(_q(x, dtype).permute(0, 3, 1, 2), _q(kern, dtype), torch.from_numpy(bias)))
    ref = ref.permute(0, 2, 3, 1).numpy()
    idx = torch.from_numpy(ops.conv_scatter_pack_index(9, 9, C, 3)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern).cuda(), idx, dtype=dtype)
    y = ops.conv2d_scatter(torch.from_numpy(x).cuda().to(dtype), wp, 3, 9, 9, bias=torch.from_numpy(bias).cuda())
    torch.cuda.synchronize()
    _close(y.cpu().numpy(), ref, 1e-3, 1e-3)


STACK_CASES = [
    # B, H, W, kh, kw, epilogue  -- all Cin = Cout = 64, stride 1: shapes the row-stacked kernel (b2_conv_stack.cu) takes
    (1, 96, 24, 3, 3, "prelu"),        # R = 3, whole tiles
    (2, 92, 21, 3, 3, "residual"),     # R = 3, ragged bottom rows and right columns
    (1, 60, 30, 3, 3, "residual"),     # R = 2 (48-row tiles would waste too much)
    (2, 47, 16, 9, 1, "prelu"),        # the folded 9x1 head conv
    (1, 48, 40, 3, 3, "gate"),         # dgrad form: ReLU gate + residual, bf16
    (1, 144, 8, 1, 3, "fp32"),         # 1x3 filter, fp32 output, several tiles per column strip
    (2, 64, 24, 5, 5, "prelu"),        # 5x5 (the reference's default kernel size): 200 KB of weights, streamed strips, R = 2
    (1, 70, 19, 5, 5, "residual"),     # 5x5 streamed, ragged edges, several tiles per CTA ring wrap
    (1, 45, 16, 5, 3, "gate"),         # 5x3 streamed (120 KB of weights), dgrad-form epilogue
]


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("case", STACK_CASES)
def test_conv2d_row_stacked(case, dtype):
    """Cout = 64 stride-1 convs run with R output rows stacked in the GEMM N dimension; same oracle, same tolerance,
    and bit-for-bit the same taps as the plain implicit GEMM (cross-checked against the CUDA-core direct kernel)."""
    from upscaler import ops
    B, H, W, kh, kw, mode = case
    C = 64
    rng = np.random.default_rng(abs(hash(case)) % 2 ** 31)
    x = _rnd((B, H, W, C), rng)
    kern = _rnd((kh, kw, C, C), rng, 1.0 / np.sqrt(kh * kw * C))
    scale = rng.uniform(0.5, 1.5, C).astype(np.float32)
    shift = _rnd((C,), rng, 0.1)
    alpha = rng.uniform(0, 0.3, C).astype(np.float32)
    res = _rnd((B, H, W, C), rng)
    gate = _rnd((B, H, W, C), rng)
    xq, kq = _q(x, dtype), _q(kern, dtype)
    ref = O.conv2d_same(xq.permute(0, 3, 1, 2), kq, None, stride=1)
    ref = ref * torch.from_numpy(scale).view(1, -1, 1, 1) + torch.from_numpy(shift).view(1, -1, 1, 1)
    kwargs = dict(scale=torch.from_numpy(scale).cuda(), shift=torch.from_numpy(shift).cuda())
    if mode == "prelu":
        ref = O.prelu(ref, torch.from_numpy(alpha))
        kwargs.update(alpha=torch.from_numpy(alpha).cuda(), act=ops.ACT_PRELU)
    if mode == "gate":
        gq = _q(gate, dtype).permute(0, 3, 1, 2)
        ref = torch.where(gq > 0, ref, ref * 0.2)
        kwargs.update(gate=torch.from_numpy(gate).cuda().to(dtype), gate_slope=0.2)
    if mode in ("residual", "gate"):
        ref = ref + _q(res, dtype).permute(0, 3, 1, 2)
        kwargs.update(residual=torch.from_numpy(res).cuda().to(dtype))
    if mode == "fp32":
        kwargs.update(out_dtype=torch.float32)
    ref = ref.permute(0, 2, 3, 1).numpy()
    idx = torch.from_numpy(ops.conv_pack_index(kh, kw, C, C)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern).cuda(), idx, dtype=dtype)
    xd = torch.from_numpy(x).cuda().to(dtype)
    y = ops.conv2d(xd, wp, C, kh, kw, **kwargs)
    yd = ops.conv2d(xd, wp, C, kh, kw, direct=True, **kwargs)
    torch.cuda.synchronize()
    assert tuple(y.shape) == ref.shape
    tol = 1e-3 if mode == "fp32" else (4e-3 if dtype == torch.float16 else 2e-2)
    _close(y.float().cpu().numpy(), ref, tol, 1e-2)
    _close(y.float().cpu().numpy(), yd.float().cpu().numpy(), tol, 1e-2)
