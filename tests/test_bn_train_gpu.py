"""-m gpu: the deterministic two-stage train-mode BatchNormalization kernels (csrc/b2_bn_train.cu) against a plain
PyTorch fp32 reference of the same op (Keras BatchNormalization, training phase: model.py:20, 23, 284, 840-884).
fp32 sums of 16-bit inputs: tolerances are fp32 round-off (1e-5 relative), moving statistics included."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
EPS, MOM = 1e-3, 0.99


def _ref_fwd(x, gamma, beta, mm, mv):
    n = x.shape[0]
    m = x.mean(0)
    v = x.var(0, unbiased=False)
    inv = 1.0 / torch.sqrt(v + EPS)
    scale = gamma * inv
    return scale, beta - m * scale, m, inv, mm * MOM + m * (1 - MOM), mv * MOM + v * (n / (n - (1 + EPS))) * (1 - MOM)


@pytest.mark.parametrize("rows,C,dt", [(16 * 96 * 96, 64, torch.bfloat16), (6 * 33 * 31, 256, torch.float16), (12, 1024, torch.bfloat16),
                                       (5, 32, torch.float16), (100000, 8, torch.bfloat16)])
def test_forward_statistics_two_stage(rows, C, dt):
    from upscaler import ops
    g = torch.Generator(device="cuda").manual_seed(rows + C)
    x = (torch.randn(rows, C, device="cuda", generator=g) * 3 + torch.linspace(-2, 2, C, device="cuda")).to(dt)
    gamma, beta = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")
    mm, mv = torch.randn(C, device="cuda"), torch.rand(C, device="cuda") + 0.5
    want = _ref_fwd(x.float(), gamma, beta, mm, mv)
    mm1, mv1 = mm.clone(), mv.clone()
    part, nparts = ops.bn_stats_partial(x)
    assert 1 <= nparts <= ops.bn_part_rows()
    aff = torch.empty(4 * C, device="cuda")
    ops.bn_fwd_finalize(part, nparts, C, rows, gamma, beta, aff, mm1, mv1)
    for got, ref, name in zip(list(aff.view(4, C)) + [mm1, mv1], want, ("scale", "shift", "mean", "invstd", "moving_mean", "moving_var")):
        err = float((got - ref).abs().max() / ref.abs().max().clamp_min(1e-6))
        assert err < 2e-5, (name, err)
    # bit-reproducible: no atomics anywhere
    part2, n2 = ops.bn_stats_partial(x)
    aff2 = torch.empty(4 * C, device="cuda")
    ops.bn_fwd_finalize(part2, n2, C, rows, gamma, beta, aff2)
    assert n2 == nparts and torch.equal(aff, aff2)
    # the summed statistics alone (NCCL fallback form), then a second finalize over them
    sums = torch.empty(2 * C, device="cuda")
    ops.bn_fwd_finalize(part, nparts, C, rows, None, None, None, sums_out=sums)
    assert torch.allclose(sums[:C], x.float().sum(0), rtol=1e-5, atol=1e-2)
    aff3 = torch.empty(4 * C, device="cuda")
    ops.bn_fwd_finalize(sums, 1, C, rows, gamma, beta, aff3)
    assert torch.equal(aff, aff3)


@pytest.mark.parametrize("act", ["none", "leaky", "prelu"])
@pytest.mark.parametrize("rows,C,dt", [(4 * 48 * 48, 64, torch.bfloat16), (6 * 8 * 8, 512, torch.float16), (7, 1024, torch.bfloat16)])
def test_backward_sums_two_stage(rows, C, dt, act):
    from upscaler import ops
    code = {"none": ops.ACT_NONE, "leaky": ops.ACT_LEAKY, "prelu": ops.ACT_PRELU}[act]
    gen = torch.Generator(device="cuda").manual_seed(rows * 3 + C)
    z = (torch.randn(rows, C, device="cuda", generator=gen) * 2 + 0.3).to(dt)
    g = torch.randn(rows, C, device="cuda", generator=gen).to(dt)
    gamma, beta = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda") * 0.3
    alpha = torch.rand(C, device="cuda") * 0.3 if act == "prelu" else None
    # reference: autograd through act(BN(z)) in fp32
    zt = z.float().requires_grad_(True)
    gm, bt = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    al = alpha.clone().requires_grad_(True) if alpha is not None else None
    m, v = zt.mean(0), zt.var(0, unbiased=False)
    u = (zt - m) / torch.sqrt(v + EPS) * gm + bt
    if act == "leaky":
        y = torch.where(u > 0, u, 0.1 * u)
    elif act == "prelu":
        y = torch.clamp(u, min=0) + al * torch.clamp(u, max=0)
    else:
        y = u
    (y * g.float()).sum().backward()
    part, nparts = ops.bn_stats_partial(z)
    aff = torch.empty(4 * C, device="cuda")
    ops.bn_fwd_finalize(part, nparts, C, rows, gamma, beta, aff)
    bpart, bn = ops.bn_bwd_partial(z, g, aff, alpha, code, 0.1)
    sums = torch.empty(2 * C, device="cuda")
    dgamma, dbeta, dalpha = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    ops.bn_bwd_finalize(bpart, bn, C, sums, dgamma, dbeta, dalpha if alpha is not None else None, 1.0)
    dz = ops.bn_bwd_apply(z, g, aff[:C], aff[C:2 * C], alpha, code, 0.1, aff[2 * C:3 * C], aff[3 * C:], sums[:C], sums[C:], rows,
                          torch.empty_like(z))
    tol = 2e-2 if dt == torch.bfloat16 else 4e-3          # dz is stored in 16 bits; the sums themselves are fp32
    rel = lambda a, b: float((a - b).norm() / b.norm().clamp_min(1e-20))
    assert rel(dbeta, bt.grad) < 1e-4 and rel(dgamma, gm.grad) < 1e-4
    if alpha is not None:
        assert rel(dalpha, al.grad) < 1e-4
    assert rel(dz.float(), zt.grad) < tol
    # a second backward pass ACCUMULATES into the gradient views and leaves its own sums intact
    sums2 = torch.empty(2 * C, device="cuda")
    ops.bn_bwd_finalize(bpart, bn, C, sums2, dgamma, dbeta, None, 0.5)
    assert torch.equal(sums, sums2)
    assert rel(dbeta, 1.5 * bt.grad) < 1e-4 and rel(dgamma, 1.5 * gm.grad) < 1e-4
    # NCCL-fallback form: raw local sums out, then finalize over the single row
    raw = torch.empty(3 * C, device="cuda")
    ops.bn_bwd_finalize(bpart, bn, C, None, raw_out=raw)
    sums3 = torch.empty(2 * C, device="cuda")
    ops.bn_bwd_finalize(raw, 1, C, sums3)
    assert torch.equal(sums, sums3)


def test_two_forward_passes_keep_their_own_statistics():
    """the real and the fake pass of the two-input discriminator model (train_gan3.py:353) run forward
    back to back through the same layers; each tape must keep the statistics of ITS batch."""
    from upscaler.networks import DiscriminatorSimple512Net
    from oracle import srgan_oracle as O
    dw = O.init_discriminator_simple_512_weights(seed=3, input_shape=(32, 32, 3))
    net = DiscriminatorSimple512Net((32, 32, 3), weights=dw)
    a = torch.randn(4, 32, 32, 3, device="cuda")
    b = torch.randn(4, 32, 32, 3, device="cuda") * 3 + 1
    _, ta = net.forward(a, True)
    affa = {k: v[1].clone() for k, v in ta.items() if isinstance(v, tuple) and len(v) == 3}
    _, tb = net.forward(b, True)
    torch.cuda.synchronize()
    assert affa
    for k, v in affa.items():
        assert torch.equal(v, ta[k][1]), k
        assert not torch.equal(ta[k][1], tb[k][1]), k
