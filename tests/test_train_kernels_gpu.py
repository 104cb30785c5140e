"""-m gpu parity of the training-step kernels (through the C ABI) against PyTorch-CPU autograd over the oracle's
TensorFlow-semantics layers (oracle/srgan_oracle.py), on the same bf16-rounded operands.

Tolerances: weight gradients are fp32 sums of bf16 products -> 2e-3 relative to the largest gradient entry; data
gradients are stored as bf16 -> 1e-2 relative + 2e-3 absolute; reductions (BN statistics, losses) 1e-3 relative.
"""
import numpy as np
import pytest
import torch

from oracle import srgan_oracle as O

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def _rnd(shape, rng, scale=1.0):
    return (rng.standard_normal(shape) * scale).astype(np.float32)


def _q(a):
    return torch.from_numpy(np.asarray(a, np.float32)).to(BF).to(torch.float32)


def _dev(a):
    return torch.from_numpy(np.asarray(a, np.float32)).cuda().to(BF)


def _close(got, ref, atol, rtol, what=""):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    err = np.abs(got - ref)
    bad = err > atol + rtol * np.abs(ref)
    assert not bad.any(), "%s: max err %.6f at %s (ref %.5f got %.5f), %d bad" % (
        what, err.max(), np.unravel_index(err.argmax(), err.shape), ref.flat[err.argmax()], got.flat[err.argmax()], bad.sum())


WGRAD_CASES = [
    # B, H, W, Cin, Cout, k, stride
    (2, 20, 33, 64, 64, 3, 1),
    (1, 17, 40, 64, 64, 5, 1),
    (1, 7, 5, 64, 64, 3, 1),        # image smaller than one 128-pixel tile (zero-padded K rows)
    (2, 24, 24, 64, 128, 3, 1),
    (1, 12, 12, 256, 256, 3, 1),    # several slot groups and channel blocks
    (2, 32, 32, 64, 128, 3, 2),     # stride 2, even input: TF pads (0,1)
    (1, 27, 15, 128, 256, 3, 2),    # stride 2, odd input: TF pads (1,1)
    (4, 1, 1, 128, 64, 1, 1),       # Dense layer as a 1x1 conv over a (B,1,1,C) tensor
    (1, 40, 40, 64, 64, 9, 1),      # 9 filter rows x folded columns (head conv form: kh=9, kw=1 is covered in the model test)
]


@pytest.mark.parametrize("case", WGRAD_CASES)
def test_conv2d_wgrad(case):
    from upscaler import ops
    B, H, W, Cin, Cout, k, s = case
    rng = np.random.default_rng(abs(hash(case)) % 2 ** 31)
    x = _rnd((B, H, W, Cin), rng)
    OH, OW = -(-H // s), -(-W // s)
    dz = _rnd((B, OH, OW, Cout), rng)
    kern = torch.zeros((k, k, Cin, Cout), requires_grad=True)
    y = O.conv2d_same(_q(x).permute(0, 3, 1, 2), kern, None, stride=s)
    (y * _q(dz).permute(0, 3, 1, 2)).sum().backward()
    ref = kern.grad.numpy()
    dw = torch.zeros((k, k, Cin, Cout), dtype=torch.float32, device="cuda")
    ops.conv2d_wgrad(_dev(x), _dev(dz), k, k, s, dw)
    ops.conv2d_wgrad(_dev(x), _dev(dz), k, k, s, dw)   # accumulates: twice the gradient
    torch.cuda.synchronize()
    _close(dw.cpu().numpy() / 2, ref, 2e-3 * np.abs(ref).max(), 1e-3, "wgrad %s" % (case,))


@pytest.mark.parametrize("k,shape", [(3, (1, 12, 20, 64, 256)), (5, (2, 9, 7, 256, 256))])
def test_conv2d_transpose_wgrad_and_dgrad(k, shape):
    """Conv2DTranspose(stride 2): wgrad = conv wgrad with the roles swapped; dgrad = stride-2 conv of the output gradient."""
    from upscaler import ops
    B, H, W, Cin, Cout = shape
    rng = np.random.default_rng(k)
    x = _rnd((B, H, W, Cin), rng)
    dout = _rnd((B, 2 * H, 2 * W, Cout), rng)
    kern_np = _rnd((k, k, Cout, Cin), rng, 1.0 / np.sqrt(k * k * Cin))
    kern = _q(kern_np).requires_grad_(True)
    xt = _q(x).permute(0, 3, 1, 2).requires_grad_(True)
    y = O.conv2d_transpose_same_s2(xt, kern, None)
    (y * _q(dout).permute(0, 3, 1, 2)).sum().backward()
    dw = torch.zeros((k, k, Cout, Cin), dtype=torch.float32, device="cuda")
    ops.conv2d_wgrad(_dev(dout), _dev(x), k, k, 2, dw)
    idx = torch.from_numpy(ops.convT_dgrad_pack_index(k, k, Cin, Cout)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern_np).cuda(), idx, dtype=BF)
    dx = ops.conv2d(_dev(dout), wp, Cin, k, k, stride=2)
    torch.cuda.synchronize()
    ref_w = kern.grad.numpy()
    _close(dw.cpu().numpy(), ref_w, 2e-3 * np.abs(ref_w).max(), 1e-3, "convT wgrad")
    ref_x = xt.grad.permute(0, 2, 3, 1).numpy()
    _close(dx.float().cpu().numpy(), ref_x, 3e-3 * np.abs(ref_x).max(), 1e-2, "convT dgrad")


@pytest.mark.parametrize("case", [(2, 20, 33, 64, 64, 3), (1, 17, 24, 64, 128, 5), (1, 12, 12, 256, 512, 3)])
def test_conv2d_dgrad_stride1_with_relu_gate(case):
    from upscaler import ops
    B, H, W, Cin, Cout, k = case
    rng = np.random.default_rng(abs(hash(case)) % 2 ** 31)
    dz = _rnd((B, H, W, Cout), rng)
    a_prev = _rnd((B, H, W, Cin), rng)           # forward output of the layer below (ReLU gate)
    res = _rnd((B, H, W, Cin), rng)
    kern_np = _rnd((k, k, Cin, Cout), rng, 1.0 / np.sqrt(k * k * Cout))
    xt = torch.zeros((B, Cin, H, W), requires_grad=True)
    y = O.conv2d_same(xt, _q(kern_np), None)
    (y * _q(dz).permute(0, 3, 1, 2)).sum().backward()
    gate = (_q(a_prev) > 0).float()
    ref = xt.grad.permute(0, 2, 3, 1) * torch.where(gate > 0, 1.0, 0.2) + _q(res)
    idx = torch.from_numpy(ops.conv_dgrad_pack_index(k, k, Cin, Cout)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern_np).cuda(), idx, dtype=BF)
    dx = ops.conv2d(_dev(dz), wp, Cin, k, k, gate=_dev(a_prev), gate_slope=0.2, residual=_dev(res))
    torch.cuda.synchronize()
    _close(dx.float().cpu().numpy(), ref.numpy(), 3e-3 * float(ref.abs().max()), 1e-2, "dgrad s1")


@pytest.mark.parametrize("case", [(2, 32, 32, 64, 128, 3), (1, 27, 15, 128, 256, 3), (3, 2, 2, 512, 512, 3), (2, 3, 3, 512, 512, 3)])
def test_conv2d_dgrad_stride2(case):
    from upscaler import ops
    B, H, W, Cin, Cout, k = case
    rng = np.random.default_rng(abs(hash(case)) % 2 ** 31)
    OH, OW = -(-H // 2), -(-W // 2)
    dz = _rnd((B, OH, OW, Cout), rng)
    kern_np = _rnd((k, k, Cin, Cout), rng, 1.0 / np.sqrt(k * k * Cout))
    xt = torch.zeros((B, Cin, H, W), requires_grad=True)
    y = O.conv2d_same(xt, _q(kern_np), None, stride=2)
    (y * _q(dz).permute(0, 3, 1, 2)).sum().backward()
    ref = xt.grad.permute(0, 2, 3, 1).numpy()
    idx = torch.from_numpy(ops.conv_s2_dgrad_pack_index(k, k, Cin, Cout)).cuda()
    wp = ops.gather_cast(torch.from_numpy(kern_np).cuda(), idx, dtype=BF)
    dx = ops.conv2d_s2_dgrad(_dev(dz), wp, Cin, k, k, (H, W))
    torch.cuda.synchronize()
    _close(dx.float().cpu().numpy(), ref, 3e-3 * np.abs(ref).max(), 1e-2, "dgrad s2")


@pytest.mark.parametrize("shape,act", [((2, 20, 24, 64), "prelu"), ((3, 9, 7, 128), "leaky"), ((5, 1, 1, 32), "leaky"),
                                       ((2, 6, 6, 512), "none")])
def test_batchnorm_train_forward_backward(shape, act):
    """colstats + bn_finalize + affine_act vs the oracle's train-mode batch_norm, and the two backward kernels vs autograd."""
    from upscaler import ops
    rng = np.random.default_rng(shape[-1])
    C = shape[-1]
    z = _rnd(shape, rng) * 1.7 + 0.3
    g = _rnd(shape, rng)
    res = _rnd(shape, rng)
    gamma = rng.uniform(0.5, 1.5, C).astype(np.float32)
    beta = _rnd((C,), rng, 0.1)
    alpha = rng.uniform(0.0, 0.3, C).astype(np.float32)
    mm = _rnd((C,), rng, 0.1)
    mv = rng.uniform(0.5, 1.5, C).astype(np.float32)
    # ---- reference
    zt = _q(z).requires_grad_(True)
    gt, bt, at = [torch.from_numpy(a).requires_grad_(True) for a in (gamma, beta, alpha)]
    perm = (0, 3, 1, 2)
    stats = []
    u = O.batch_norm(zt.permute(*perm), gt, bt, None, None, True, stats)
    if act == "prelu":
        y = O.prelu(u, at)
    elif act == "leaky":
        y = O.leaky(u, 0.1)
    else:
        y = u
    y = y + _q(res).permute(*perm)
    (y * _q(g).permute(*perm)).sum().backward()
    bm, bv, n = stats[0]
    ref_mm, ref_mv = O.update_moving(torch.from_numpy(mm), torch.from_numpy(mv), bm, bv, n)
    # ---- device
    f = lambda a: torch.from_numpy(a).cuda()
    zd, gd, rd = _dev(z), _dev(g), _dev(res)
    ssum, ssq = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    scale, shift, mean, invstd = [torch.empty(C, device="cuda") for _ in range(4)]
    mmd, mvd = f(mm.copy()), f(mv.copy())
    ops.colstats(zd, ssum, ssq)
    rows = zd.numel() // C
    ops.bn_finalize(ssum, ssq, rows, f(gamma), f(beta), scale, shift, mean, invstd, mmd, mvd)
    code = {"prelu": ops.ACT_PRELU, "leaky": ops.ACT_LEAKY, "none": ops.ACT_NONE}[act]
    al = f(alpha) if act == "prelu" else None
    yd = ops.affine_act(zd, torch.empty_like(zd), scale, shift, al, code, 0.1, rd)
    s1, s2, da = [torch.zeros(C, device="cuda") for _ in range(3)]
    ops.bn_bwd_reduce(zd, gd, scale, shift, al, code, 0.1, mean, invstd, s1, s2, da if act == "prelu" else None)
    dzd = ops.bn_bwd_apply(zd, gd, scale, shift, al, code, 0.1, mean, invstd, s1, s2, rows, torch.empty_like(zd))
    torch.cuda.synchronize()
    _close(mean.cpu().numpy(), bm.detach().numpy(), 1e-4, 1e-3, "batch mean")
    _close(mmd.cpu().numpy(), ref_mm.detach().numpy(), 1e-4, 1e-3, "moving mean")
    _close(mvd.cpu().numpy(), ref_mv.detach().numpy(), 1e-4, 1e-3, "moving var")
    _close(yd.float().cpu().numpy(), y.detach().permute(0, 2, 3, 1).numpy(), 2e-2, 1e-2, "bn+act forward")
    _close(s1.cpu().numpy(), bt.grad.numpy(), 2e-3 * float(bt.grad.abs().max()) + 1e-4, 2e-3, "dbeta")
    _close(s2.cpu().numpy(), gt.grad.numpy(), 2e-3 * float(gt.grad.abs().max()) + 1e-4, 2e-3, "dgamma")
    if act == "prelu":
        _close(da.cpu().numpy(), at.grad.numpy(), 2e-3 * float(at.grad.abs().max()) + 1e-4, 2e-3, "dalpha")
    refdz = zt.grad.numpy()
    _close(dzd.float().cpu().numpy(), refdz, 1e-2 * np.abs(refdz).max(), 1e-2, "dz")


def test_bn_fold_matches_inference_batch_norm():
    from upscaler import ops
    rng = np.random.default_rng(3)
    C = 64
    gamma, beta, mean, var, bias = (rng.uniform(0.5, 1.5, C), rng.normal(0, .1, C), rng.normal(0, .1, C), rng.uniform(.5, 1.5, C),
                                    rng.normal(0, .1, C))
    f = lambda a: torch.from_numpy(a.astype(np.float32)).cuda()
    scale, shift = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    ops.bn_fold(f(gamma), f(beta), f(mean), f(var), f(bias), scale, shift)
    sc = gamma / np.sqrt(var + 1e-3)
    _close(scale.cpu().numpy(), sc, 1e-5, 1e-5)
    _close(shift.cpu().numpy(), beta + (bias - mean) * sc, 1e-5, 1e-5)


@pytest.mark.parametrize("shape", [(2, 8, 12, 64), (1, 7, 9, 128)])
def test_maxpool_forward_backward(shape):
    from upscaler import ops
    rng = np.random.default_rng(shape[1])
    B, H, W, C = shape
    x = np.maximum(_rnd(shape, rng), 0)            # post-ReLU input, ties at zero
    dy = _rnd((B, H // 2, W // 2, C), rng)
    xt = _q(x).permute(0, 3, 1, 2).requires_grad_(True)
    y = torch.nn.functional.max_pool2d(xt, 2, 2)
    (y * _q(dy).permute(0, 3, 1, 2)).sum().backward()
    xd = _dev(x)
    yd = ops.maxpool2x2(xd)
    dxd = ops.maxpool2x2_bwd(xd, yd, _dev(dy), relu_gate=True)
    torch.cuda.synchronize()
    _close(yd.float().cpu().numpy(), y.detach().permute(0, 2, 3, 1).numpy(), 0, 0, "maxpool")
    ref = (xt.grad * (xt.detach() > 0)).permute(0, 2, 3, 1).numpy()      # ties only happen at 0 where the ReLU gate kills them
    _close(dxd.float().cpu().numpy(), ref, 0, 0, "maxpool bwd")


@pytest.mark.parametrize("mae", [False, True])
def test_loss_pair_and_tanh_bwd(mae):
    from upscaler import ops
    rng = np.random.default_rng(5)
    a, b = _rnd((2, 12, 12, 512), rng), _rnd((2, 12, 12, 512), rng)
    at = _q(a).requires_grad_(True)
    d = at - _q(b)
    l = (d.abs() if mae else d * d).mean() * 0.7
    l.backward()
    loss = torch.zeros(1, device="cuda")
    grad = torch.empty((2, 12, 12, 512), dtype=BF, device="cuda")
    ops.loss_pair(_dev(a), _dev(b), loss, mae=mae, weight=0.7, gscale=0.7, grad=grad)
    y = np.tanh(_rnd((2, 8, 8, 3), rng))
    g1, g2 = _rnd((2, 8, 8, 3), rng), _rnd((2, 8, 8, 3), rng)
    f = lambda v: torch.from_numpy(v).cuda()
    dz = ops.tanh_bwd(f(y), f(g1), torch.empty((2, 8, 8, 3), device="cuda"), g2=f(g2))
    torch.cuda.synchronize()
    _close(loss.cpu().numpy(), [float(l)], 1e-6, 1e-3, "loss")
    _close(grad.float().cpu().numpy(), at.grad.numpy(), 1e-9, 1e-2, "loss grad")
    _close(dz.cpu().numpy(), (g1 + g2) * (1 - y * y), 1e-6, 1e-5, "tanh bwd")


def test_adam_matches_keras_adam():
    from upscaler import ops
    rng = np.random.default_rng(11)
    n = 1003
    p0 = _rnd((n,), rng)
    opt = O.KerasAdam()
    pt = torch.from_numpy(p0.copy())
    p = torch.from_numpy(p0.copy()).cuda()
    m, v = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    for t in range(1, 4):
        g = _rnd((n,), rng, 10.0 ** (-t))
        pt.grad = torch.from_numpy(g.copy())
        opt.apply([("p", pt)])
        lr_t = 1e-3 * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        ops.adam(p, torch.from_numpy(g * 4).cuda(), m, v, float(lr_t), 0.9, 0.999, 1e-7, gscale=0.25)
    torch.cuda.synchronize()
    _close(p.cpu().numpy(), pt.numpy(), 1e-6, 1e-5, "adam")


VALID_CASES = [
    # B, H, W, Cin, Cout, k, stride       (make_discriminator_sparse_512, model.py:967-989: 5x5 'valid', strides 1 / 3)
    (2, 40, 37, 64, 128, 5, 3),
    (1, 168, 100, 64, 128, 5, 3),         # wide rows: the strided TMA box limits the tile width
    (3, 17, 17, 128, 256, 5, 3),
    (4, 5, 5, 256, 256, 5, 3),            # 5x5 -> 1x1
    (2, 20, 23, 64, 64, 5, 1),
    (2, 19, 22, 64, 128, 3, 2),
]


@pytest.mark.parametrize("case", VALID_CASES)
def test_conv2d_valid_forward_dgrad_wgrad(case):
    """b2_conv2d_fprop_pad / b2_conv2d_strided_dgrad / b2_conv2d_wgrad_pad against autograd of a plain fp32
    conv2d(padding=0) on the same bf16-rounded operands."""
    import torch.nn.functional as F
    from upscaler import ops
    B, H, W, Cin, Cout, k, s = case
    rng = np.random.default_rng(abs(hash(case)) % 2 ** 31)
    OH, OW = (H - k) // s + 1, (W - k) // s + 1
    x = _rnd((B, H, W, Cin), rng)
    dz = _rnd((B, OH, OW, Cout), rng)
    kern_np = _rnd((k, k, Cin, Cout), rng, 1.0 / np.sqrt(k * k * Cin))
    bias = _rnd((Cout,), rng, 0.1)
    xt = _q(x).permute(0, 3, 1, 2).clone().requires_grad_(True)
    kt = _q(kern_np).requires_grad_(True)
    y = F.conv2d(xt, kt.permute(3, 2, 0, 1), torch.from_numpy(bias), stride=s)
    assert tuple(y.shape) == (B, Cout, OH, OW)
    (y * _q(dz).permute(0, 3, 1, 2)).sum().backward()
    kd = torch.from_numpy(kern_np).cuda()
    wp = ops.gather_cast(kd, torch.from_numpy(ops.conv_pack_index(k, k, Cin, Cout)).cuda(), dtype=BF)
    got = ops.conv2d_pad(_dev(x), wp, Cout, k, k, s, (0, 0), (OH, OW), shift=torch.from_numpy(bias).cuda())
    ref = y.detach().permute(0, 2, 3, 1).numpy()
    _close(got.float().cpu().numpy(), ref, 4e-3 * np.abs(ref).max(), 1e-2, "valid fprop")
    ref_dx = xt.grad.permute(0, 2, 3, 1).numpy()
    if s > 1:
        wt = ops.gather_cast(kd, torch.from_numpy(ops.conv_s2_dgrad_pack_index(k, k, Cin, Cout)).cuda(), dtype=BF)
        dx = ops.conv2d_strided_dgrad(_dev(dz), wt, Cin, k, k, s, (0, 0), (H, W))
    else:
        wt = ops.gather_cast(kd, torch.from_numpy(ops.conv_dgrad_pack_index(k, k, Cin, Cout)).cuda(), dtype=BF)
        dx = ops.conv2d_pad(_dev(dz), wt, Cin, k, k, 1, (k - 1, k - 1), (H, W))
    _close(dx.float().cpu().numpy(), ref_dx, 4e-3 * np.abs(ref_dx).max(), 1e-2, "valid dgrad")
    dw = torch.zeros((k, k, Cin, Cout), device="cuda")
    ops.conv2d_wgrad_pad(_dev(x), _dev(dz), k, k, s, (0, 0), dw)
    torch.cuda.synchronize()
    ref_dw = kt.grad.numpy()
    _close(dw.cpu().numpy(), ref_dw, 2e-3 * np.abs(ref_dw).max(), 2e-3, "valid wgrad")
