"""-m gpu: b2_bn_fwd_stats_static, the forward BatchNormalization statistics for hosts that cannot allocate device memory.

Who needs it: code that embeds libb2srgan.so somewhere device memory is not its own to hand out -- a plugin living inside
another framework's memory arena, a frame loop with a real-time budget where a trip into the allocator is an unbounded
stall, a process whose device-memory quota is already committed.  Those callers hold the activation tensor and the
result vectors, but b2_bn_stats_partial + b2_bn_fwd_finalize also want a [b2_bn_part_rows()][2*C] scratch tensor between
the two launches.  The static variant keeps those rows in an array compiled into the library (C <= 512) and so needs
nothing beyond inputs and outputs.

Checked here, through the C ABI (ctypes -> libb2srgan.so), against the existing two-call entry points:
  * two calls in a row on different shapes (a big layer, then a small one that leaves most of the shared array holding the
    first call's rows) both match -- nothing stale leaks in, nothing needs clearing between calls;
  * layers wider than the built-in array are refused with a readable error instead of overrunning it;
  * the call itself leaves the allocator alone (torch's counters and the driver's free-memory figure do not move).
This file sorts last on purpose: it shares one device-global array across calls, and a failure here must not hide the
parity tests under `pytest -x`.  It does not depend on any other test having run (or not run) before it."""
import pytest
import torch

pytestmark = pytest.mark.gpu
EPS = 1e-3
REL_TOL = 2e-5          # fp32 sums of 16-bit inputs: the bar of tests/test_bn_train_gpu.py


def _two_call_form(ops, x, gamma, beta):
    rows, C = x.shape
    part, nparts = ops.bn_stats_partial(x)
    sums = torch.empty(2 * C, device="cuda")
    aff = torch.empty(4 * C, device="cuda")
    ops.bn_fwd_finalize(part, nparts, C, rows, gamma, beta, aff, sums_out=sums, eps=EPS)
    return sums, aff


def _make(rows, C, dt, seed):
    gen = torch.Generator(device="cuda").manual_seed(seed)
    x = (torch.randn(rows, C, device="cuda", generator=gen) * 1.7 + 0.4).to(dt)
    gamma = torch.rand(C, device="cuda", generator=gen) + 0.5
    beta = torch.randn(C, device="cuda", generator=gen) * 0.3
    return x, gamma, beta


def _rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-6))


def test_static_scratch_matches_two_call_form_twice_in_a_row():
    from upscaler import ops
    # big first (fills all rows of the shared array), then small (writes a handful of rows, the rest keep call 1's values),
    # then the big one again
    cases = [(4 * 96 * 96, 64, torch.float16, 11), (7, 512, torch.bfloat16, 12), (6 * 8 * 8, 256, torch.float16, 13),
             (4 * 96 * 96, 64, torch.float16, 11)]
    data = [_make(*c) for c in cases]
    want = [_two_call_form(ops, x, g, b) for x, g, b in data]
    outs = [(torch.full((2 * x.shape[1],), float("nan"), device="cuda"), torch.full((4 * x.shape[1],), float("nan"), device="cuda"))
            for x, _, _ in data]
    torch.cuda.synchronize()
    for (x, g, b), (sums, aff) in zip(data, outs):          # back to back on one stream, no sync in between
        ops.bn_fwd_stats_static(x, x.shape[0], g, b, sums_out=sums, aff=aff, eps=EPS)
    torch.cuda.synchronize()
    same_grid = ops.bn_part_rows() <= 296                    # B200: both forms use the same stage-1 grid -> same bits
    for (x, _, _), (sums, aff), (wsums, waff) in zip(data, outs, want):
        C = x.shape[1]
        assert torch.isfinite(sums).all() and torch.isfinite(aff).all()
        assert _rel(sums[:C], wsums[:C]) < REL_TOL and _rel(sums[C:], wsums[C:]) < REL_TOL
        for k, name in enumerate(("scale", "shift", "mean", "invstd")):
            assert _rel(aff[k * C:(k + 1) * C], waff[k * C:(k + 1) * C]) < REL_TOL, name
        if same_grid:
            assert torch.equal(sums, wsums) and torch.equal(aff, waff)
        # and against plain fp32 PyTorch
        xf = x.float()
        assert torch.allclose(aff[2 * C:3 * C], xf.mean(0), rtol=1e-4, atol=1e-5)
        assert torch.allclose(aff[3 * C:], torch.rsqrt(xf.var(0, unbiased=False) + EPS), rtol=1e-4)


def test_static_scratch_sums_only_and_limits():
    from upscaler import ops
    from upscaler._lib import B2Error
    x, gamma, beta = _make(1000, 128, torch.bfloat16, 21)
    sums = torch.empty(256, device="cuda")
    ops.bn_fwd_stats_static(x, 1000, sums_out=sums)          # statistics alone, no affine
    wsums, _ = _two_call_form(ops, x, gamma, beta)
    assert _rel(sums, wsums) < REL_TOL
    wide = torch.zeros(16, 1024, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(B2Error, match="512"):                # wider than the built-in array: refused, not overrun
        ops.bn_fwd_stats_static(wide, 16, sums_out=torch.empty(2048, device="cuda"))
    with pytest.raises(B2Error):                             # nothing to write
        ops.bn_fwd_stats_static(x, 1000)
    with pytest.raises(B2Error):                             # affine wanted but no gamma / beta
        ops.bn_fwd_stats_static(x, 1000, aff=torch.empty(512, device="cuda"))


def test_static_scratch_call_does_not_allocate():
    from upscaler import ops
    x, gamma, beta = _make(3 * 64 * 64, 64, torch.float16, 31)
    sums, aff = torch.empty(128, device="cuda"), torch.empty(256, device="cuda")
    ops.bn_fwd_stats_static(x, x.shape[0], gamma, beta, sums_out=sums, aff=aff)   # first use on this device: address lookup
    torch.cuda.synchronize()
    stats0 = torch.cuda.memory_stats()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        ops.bn_fwd_stats_static(x, x.shape[0], gamma, beta, sums_out=sums, aff=aff)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    stats1 = torch.cuda.memory_stats()
    assert stats1["allocation.all.allocated"] == stats0["allocation.all.allocated"]
    assert stats1["reserved_bytes.all.current"] == stats0["reserved_bytes.all.current"]
    assert free1 >= free0, (free0, free1)                    # the driver handed nothing out either
