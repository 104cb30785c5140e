"""-m gpu parity of the discriminator and of the adversarial training steps (BASELINE.json config 4: train_gan.py
make_and_compile_gan, train_gan3.py make_and_compile_gan2) against the fp32 CPU oracle (GanTrainer / Gan2Trainer).
Tolerances as in test_training_gpu.py (bf16 storage: forward within 1-3 %, gradients by norm and direction)."""
import numpy as np
import pytest
import torch

from oracle import srgan_oracle as O

pytestmark = pytest.mark.gpu
HR = 64
# adversarial losses are (differences of) batch means of critic scores of magnitude O(1); with bf16 activation storage (the
# production type: the critic's range needs it) each mean carries a few % of rounding error through eleven batch-statistics
# BatchNorms over 8 samples.  Measured over repeated runs: offsets of 0.021-0.033 with +-0.006 of atomic-order noise
# (profiles/r01_s46_gan2_repeat.log); with fp16 storage the offset is 0.005-0.007 (test_gan2_disc_loss_fp16_storage).
ADV_ATOL = 5e-2


def _rel(got, ref):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    return float(np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-30))


def _data(B, lr, seed=0):
    rng = np.random.default_rng(seed)
    hr = rng.uniform(-1, 1, (B, 4 * lr, 4 * lr, 3)).astype(np.float32)
    hr = (hr + np.roll(hr, 1, 1) + np.roll(hr, 1, 2) + np.roll(hr, (1, 1), (1, 2))) / 4
    x = hr.reshape(B, lr, 4, lr, 4, 3).mean(axis=(2, 4)).astype(np.float32)
    return x, hr.astype(np.float32)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(16, 64, 64), (12, 40, 56)])
def test_discriminator_forward_backward(shape, dtype):
    """The deep blocks normalise over very few values (n x 1 x 1) and LeakyReLU(0.1) gates change a gradient element
    tenfold when they flip, so this network amplifies storage rounding strongly.  fp16 storage (11-bit significand)
    pins the implementation: norms within 5 %, cos >= 0.98; bf16 storage (the production type, needed for the critic's
    range) is held to norms within 20 %, cos >= 0.9."""
    from upscaler.networks import DiscriminatorSimple512Net
    tol_n, tol_c = (0.06, 0.97) if dtype == torch.float16 else (0.2, 0.9)   # block_3 BN at 12x40x56 sits at cos 0.981-0.983 run to run (atomic summation order)
    n, H, W = shape
    dw = O.init_discriminator_simple_512_weights(seed=3, input_shape=(H, W, 3))
    rng = np.random.default_rng(n)
    x = rng.uniform(-1, 1, (n, H, W, 3)).astype(np.float32)
    lab = np.where(np.arange(n) % 2 == 0, 1.0, -1.0).astype(np.float32)
    w = O.to_torch(dw, requires_grad=True)
    xt = torch.from_numpy(x).requires_grad_(True)
    with torch.no_grad():
        inf_ref = O.discriminator_simple_512_forward(w, x, "none", False).numpy()
    out = O.discriminator_simple_512_forward(w, xt, "none", True, [])
    loss = O.wasserstein_loss(torch.from_numpy(lab).reshape(-1, 1), out)
    loss.backward()
    net = DiscriminatorSimple512Net((H, W, 3), weights=dw, dtype=dtype)
    xd = torch.from_numpy(x).cuda()
    inf = net.forward(xd, False)[0]
    scores, tape = net.forward(xd, True)
    from upscaler import ops
    lb = torch.zeros(1, device="cuda")
    ds = torch.empty_like(scores)
    ops.mean_prod(scores, torch.from_numpy(lab).cuda(), lb, 1.0, 1.0, ds)
    net.ensure_grads().zero_()
    dimg = net.backward(tape, ds, True, True)
    torch.cuda.synchronize()
    print("D %s: inference scores rel err %.4f, train scores rel err %.4f, loss %.5f vs %.5f" % (
        shape, _rel(inf.cpu().numpy(), inf_ref), _rel(scores.cpu().numpy(), out.detach().numpy()), float(lb), float(loss)))
    assert _rel(inf.cpu().numpy(), inf_ref) < (5e-3 if dtype == torch.float16 else 3e-2)
    assert _rel(scores.cpu().numpy(), out.detach().numpy()) < (1e-2 if dtype == torch.float16 else 8e-2)
    bad = []
    for name, kind, cfg in net.specs:
        for j, gt in enumerate(net.grads[name]):
            ref = w[name][j].grad
            if ref is None or (kind in ("conv", "dense") and j == 1 and name != "discriminator/final/Dense_3"):
                continue      # biases in front of a train-mode BatchNorm: analytically zero gradient
            ref = ref.numpy()
            got = gt.cpu().numpy().reshape(ref.shape)
            if np.linalg.norm(ref) < 1e-6:      # e.g. Dense_3 bias under balanced +1/-1 labels
                assert np.linalg.norm(got) < 1e-4
                continue
            nr = np.linalg.norm(got) / max(np.linalg.norm(ref), 1e-30)
            cos = float(np.vdot(got, ref) / max(np.linalg.norm(got) * np.linalg.norm(ref), 1e-30))
            print("   grad %-40s[%d] norm ratio %.4f cos %.4f (|ref| %.2e)" % (name, j, nr, cos, np.linalg.norm(ref)))
            if abs(nr - 1) > tol_n or cos < tol_c:
                bad.append((name, j, nr, cos))
    gi = xt.grad.numpy()
    nr, cos = np.linalg.norm(dimg.cpu().numpy()) / np.linalg.norm(gi), float(np.vdot(dimg.cpu().numpy(), gi) / (np.linalg.norm(dimg.cpu().numpy()) * np.linalg.norm(gi)))
    print("   input gradient norm ratio %.4f cos %.4f" % (nr, cos))
    assert not bad, bad
    assert abs(nr - 1) < tol_n and cos > tol_c


@pytest.mark.parametrize("act", ["sigmoid", "log-sigm", "tanh", "bi-log"])
def test_score_activation_kernel(act):
    """b2_score_act (discriminator output layer, model.py:885-892; train_gan3.py -da) against autograd of the oracle's
    restatement: fp32, so tight."""
    from upscaler import ops
    z = np.concatenate((np.linspace(-12, 12, 97), [0.0, 1e-4, -1e-4, 40.0, -40.0])).astype(np.float32).reshape(-1, 1)
    dy = np.random.default_rng(1).normal(size=z.shape).astype(np.float32)
    zt = torch.from_numpy(z).requires_grad_(True)
    ref = O.loss_activation(act, zt)
    ref.backward(torch.from_numpy(dy))
    zd = torch.from_numpy(z).cuda()
    got = ops.score_act(zd, act).cpu().numpy()
    gback = ops.score_act(zd, act, dy=torch.from_numpy(dy).cuda()).cpu().numpy()
    # float32 log(sigmoid(x)) loses everything below -17 or so in the oracle's two-step form; compare where it is exact
    ok = np.abs(z) <= 12 if act == "log-sigm" else np.ones_like(z, bool)
    assert np.allclose(got[ok], ref.detach().numpy()[ok], rtol=2e-5, atol=2e-6)
    assert np.allclose(gback[ok], zt.grad.numpy()[ok], rtol=2e-5, atol=2e-6)
    assert np.isfinite(got).all() and np.isfinite(gback).all()


@pytest.mark.parametrize("act", ["sigmoid", "log-sigm", "tanh", "bi-log"])
def test_discriminator_output_activation(act):
    """make_discriminator_simple_512(..., activation=act): scores and the image gradient through the output layer."""
    from upscaler import ops
    from upscaler.model import make_discriminator_simple_512
    n, H, W = 16, 64, 64
    dw = O.init_discriminator_simple_512_weights(seed=5, input_shape=(H, W, 3))
    x = np.random.default_rng(2).uniform(-1, 1, (n, H, W, 3)).astype(np.float32)
    lab = np.where(np.arange(n) % 2 == 0, 1.0, -1.0).astype(np.float32)
    w = O.to_torch(dw, requires_grad=True)
    xt = torch.from_numpy(x).requires_grad_(True)
    with torch.no_grad():
        inf_ref = O.discriminator_simple_512_forward(w, x, act, False).numpy()
    out = O.discriminator_simple_512_forward(w, xt, act, True, [])
    O.wasserstein_loss(torch.from_numpy(lab).reshape(-1, 1), out).backward()
    D = make_discriminator_simple_512((H, W, 3), activation=act)
    D.net.set_weights(dw)
    assert np.abs(D.predict(x) - inf_ref).max() < 3e-2 * max(1.0, np.abs(inf_ref).max())
    xd = torch.from_numpy(x).cuda()
    scores, tape = D.net.forward(xd, True)
    ds = torch.empty_like(scores)
    ops.mean_prod(scores, torch.from_numpy(lab).cuda(), torch.zeros(1, device="cuda"), 1.0, 1.0, ds)
    D.net.ensure_grads().zero_()
    dimg = D.net.backward(tape, ds, True, True).cpu().numpy()
    ref_s = out.detach().numpy()
    print("%s: train scores max abs err %.4f (|ref| max %.3f)" % (act, np.abs(scores.cpu().numpy() - ref_s).max(), np.abs(ref_s).max()))
    assert np.abs(scores.cpu().numpy() - ref_s).max() < 8e-2 * max(1.0, np.abs(ref_s).max())
    gi = xt.grad.numpy()
    nr = np.linalg.norm(dimg) / np.linalg.norm(gi)
    cos = float(np.vdot(dimg, gi) / (np.linalg.norm(dimg) * np.linalg.norm(gi)))
    print("   input gradient norm ratio %.4f cos %.4f" % (nr, cos))
    assert abs(nr - 1) < 0.2 and cos > 0.9
    gw = D.net.grads["discriminator/final/Dense_3"][0].cpu().numpy().ravel()
    rw = w["discriminator/final/Dense_3"][0].grad.numpy().ravel()
    assert abs(np.linalg.norm(gw) / np.linalg.norm(rw) - 1) < 0.2


def _gan_models(kind2=False, family="relativistic", act="log-sigm", disc="simple", HR=HR, dact="none", ddtype=None, rbn=1,
                dlw=1e-3):
    from upscaler.model import (RelativisticLosses, VGG_MSE_LOSS, WassersteinLosses, make_and_compile_gan,
                                make_and_compile_gan2, make_discriminator_simple_512, make_discriminator_sparse_512,
                                make_upscaler_orig, wasserstein_loss)
    from upscaler.keras_compat import Adam
    gw = O.init_upscaler_orig_weights(seed=21, kernel_size=3, res_block_num=rbn)
    if disc == "sparse":
        dw = O.init_discriminator_sparse_512_weights(seed=22, input_shape=(HR, HR, 3))
        make_d, dfwd = make_discriminator_sparse_512, O.discriminator_sparse_512_forward
    else:
        dw = O.init_discriminator_simple_512_weights(seed=22, input_shape=(HR, HR, 3))
        make_d, dfwd = make_discriminator_simple_512, O.discriminator_simple_512_forward
    vw = O.init_vgg19_weights()
    G = make_upscaler_orig((HR, HR, 3), kernel_size=3, res_block_num=rbn)
    G.net.set_weights(gw)
    D = make_d((HR, HR, 3), activation=dact)
    if ddtype is not None:
        from upscaler.keras_compat import Model
        D = Model(net=type(D.net)((HR, HR, 3), activation=dact, dtype=ddtype), input_shape=(HR, HR, 3), name="model_2")
    D.net.set_weights(dw)
    content = VGG_MSE_LOSS((HR, HR, 3), 1e-3, vgg19=vw).loss
    # dlw = 1e-3 by default: large enough for the adversarial term to matter in the comparison (reference default 1e-10 / 1e-5)
    if not kind2:
        models = make_and_compile_gan(G, D, (HR // 4, HR // 4, 3), (HR, HR, 3), content, 1, wasserstein_loss, dlw, optimizer=Adam())
        ref = O.GanTrainer(gw, dw, vw, "vgg-mse", 1e-3, 1.0, dlw, res_block_num=rbn, disc_activation=dact, disc_forward=dfwd)
    else:
        mk = (lambda: WassersteinLosses()) if family == "wasserstein" else (lambda: RelativisticLosses(loss_activation=act))
        models = make_and_compile_gan2(G, D, (HR // 4, HR // 4, 3), (HR, HR, 3), content, 1, mk, dlw, optimizer=Adam())
        ref = O.Gan2Trainer(gw, dw, vw, "vgg-mse", 1e-3, 1.0, dlw, res_block_num=rbn, family=family, loss_act=act,
                            disc_activation=dact, disc_forward=dfwd)
    return G, D, models, ref


def test_gan_step_train_gan_py():
    """train_gan.py:305-317: predict -> disc_train.train_on_batch -> gan_train.train_on_batch, labels +1 / -1."""
    G, D, (gen_train, disc_train, gan_train), ref = _gan_models()
    x, hr = _data(8, HR // 4, seed=5)
    n = x.shape[0]
    gen = gen_train.predict(x)
    assert np.abs(gen - ref.predict(x)).max() < 2e-2
    d_in = np.concatenate((hr, gen), axis=0)
    d_y = np.concatenate((np.ones(n), -np.ones(n)), axis=0)
    ld = disc_train.train_on_batch(d_in, d_y)
    ld_ref = ref.disc_train_on_batch(d_in, d_y)
    print("disc loss %.5f vs oracle %.5f" % (ld, ld_ref))
    assert abs(ld - ld_ref) <= 3e-2 * abs(ld_ref) + ADV_ATOL
    # same discriminator on both sides before the generator step (its first Adam step is sign-like, see test_training_gpu)
    D.net.set_weights({k: [t.detach().numpy() for t in ts] for k, ts in ref.d.items()})
    lg = gan_train.train_on_batch(x, [hr, np.ones(n)])
    lg_ref = ref.gan_train_on_batch(x, hr, np.ones(n, np.float32))
    print("gan losses %s vs oracle %s" % (np.round(lg, 5), np.round(lg_ref, 5)))
    assert len(lg) == 3
    assert abs(lg[1] - lg_ref[1]) <= 1e-2 * abs(lg_ref[1])
    assert abs(lg[2] - lg_ref[2]) <= 3e-2 * abs(lg_ref[2]) + ADV_ATOL
    assert abs(lg[0] - lg_ref[0]) <= 1e-2 * abs(lg_ref[0])
    # the frozen discriminator's weights did not move in the GAN step, its moving statistics did (Keras semantics)
    k = "discriminator/block_3/Conv2d"
    assert np.array_equal(D.net.weights[k][0].cpu().numpy(), ref.d[k][0].detach().numpy())
    # shared Adam `iterations`: two updates happened on the one default-style optimizer object
    assert gan_train._compiled.opt.iterations == 2 and gan_train._compiled.opt is disc_train._compiled.opt


@pytest.mark.parametrize("family,act", [("wasserstein", "none"), ("relativistic", "log-sigm"), ("relativistic", "bi-log")])
def test_gan2_step_train_gan3_py(family, act):
    """train_gan3.py:339-354: two-input discriminator model and generator loss built from D(real), D(fake)."""
    G, D, (gen_train, disc_train, gan_train), ref = _gan_models(True, family, act)
    x, hr = _data(8, HR // 4, seed=6)
    n = x.shape[0]
    gen = gen_train.predict(x)
    ld = disc_train.train_on_batch([hr, gen], -np.ones(n))
    ld_ref = ref.disc_train_on_batch2(hr, gen)
    print("%s/%s disc loss %.5f vs oracle %.5f" % (family, act, ld, ld_ref))
    assert abs(ld - ld_ref) <= 3e-2 * abs(ld_ref) + ADV_ATOL
    D.net.set_weights({k: [t.detach().numpy() for t in ts] for k, ts in ref.d.items()})
    lg = gan_train.train_on_batch([x, hr], [hr, -np.ones(n)])
    lg_ref = ref.gan_train_on_batch2(x, hr, hr)
    print("%s/%s gan losses %s vs oracle %s" % (family, act, np.round(lg, 5), np.round(lg_ref, 5)))
    assert abs(lg[1] - lg_ref[1]) <= 1e-2 * abs(lg_ref[1])
    assert abs(lg[2] - lg_ref[2]) <= 3e-2 * abs(lg_ref[2]) + ADV_ATOL


def test_discriminator_thin_512_forward_backward():
    """make_discriminator_thin_512 (model.py:901-960): same kernels, 128-wide blocks; scores and gradients vs the oracle."""
    from upscaler.networks import DiscriminatorThin512Net
    from upscaler import ops
    n, H, W = 8, 64, 64
    dw = O.init_discriminator_simple_512_weights(seed=5, input_shape=(H, W, 3), filters=O.DISC_THIN_FILTERS)
    x = np.random.default_rng(2).uniform(-1, 1, (n, H, W, 3)).astype(np.float32)
    lab = np.where(np.arange(n) % 2 == 0, 1.0, -1.0).astype(np.float32)
    w = O.to_torch(dw, requires_grad=True)
    with torch.no_grad():
        inf_ref = O.discriminator_simple_512_forward(w, x, "none", False).numpy()
    out = O.discriminator_simple_512_forward(w, torch.from_numpy(x), "none", True, [])
    O.wasserstein_loss(torch.from_numpy(lab).reshape(-1, 1), out).backward()
    net = DiscriminatorThin512Net((H, W, 3), weights=dw, dtype=torch.float16)
    assert sum(int(np.prod(t.shape)) for v in dw.values() for t in v) == sum(int(t.numel()) for v in net.weights.values() for t in v)
    xd = torch.from_numpy(x).cuda()
    inf = net.forward(xd, False)[0]
    scores, tape = net.forward(xd, True)
    lb = torch.zeros(1, device="cuda")
    ds = torch.empty_like(scores)
    ops.mean_prod(scores, torch.from_numpy(lab).cuda(), lb, 1.0, 1.0, ds)
    net.ensure_grads().zero_()
    net.backward(tape, ds, True, False)
    torch.cuda.synchronize()
    assert _rel(inf.cpu().numpy(), inf_ref) < 5e-3
    assert _rel(scores.cpu().numpy(), out.detach().numpy()) < 2e-2     # batch-statistics BN over 8 samples; seen 0.008-0.0103 run to run
    for name in ("discriminator/block_2/Conv2d", "discriminator/block_5/Conv2d", "discriminator/final/Dense_1"):
        ref = w[name][0].grad.numpy()
        got = net.grads[name][0].cpu().numpy().reshape(ref.shape)
        cos = float(np.vdot(got, ref) / (np.linalg.norm(got) * np.linalg.norm(ref)))
        print("thin_512 grad %s cos %.4f norm ratio %.4f" % (name, cos, np.linalg.norm(got) / np.linalg.norm(ref)))
        assert cos > 0.97 and abs(np.linalg.norm(got) / np.linalg.norm(ref) - 1) < 0.06


def test_discriminator_sparse_512_forward_backward():
    """make_discriminator_sparse_512 (model.py:964-1012): 5x5 'valid' blocks with strides 1,3,3,3,3,3 at the smallest
    input it accepts in one dimension (489) and a ragged other one; scores, weight gradients and the image gradient."""
    from upscaler.networks import DiscriminatorSparse512Net
    from upscaler import ops
    n, H, W = 6, 489, 500
    dw = O.init_discriminator_sparse_512_weights(seed=7, input_shape=(H, W, 3))
    x = np.random.default_rng(3).uniform(-1, 1, (n, H, W, 3)).astype(np.float32)
    lab = np.where(np.arange(n) % 2 == 0, 1.0, -1.0).astype(np.float32)
    w = O.to_torch(dw, requires_grad=True)
    xt = torch.from_numpy(x).requires_grad_(True)
    with torch.no_grad():
        inf_ref = O.discriminator_sparse_512_forward(w, x, "none", False).numpy()
    out = O.discriminator_sparse_512_forward(w, xt, "none", True, [])
    O.wasserstein_loss(torch.from_numpy(lab).reshape(-1, 1), out).backward()
    net = DiscriminatorSparse512Net((H, W, 3), weights=dw, dtype=torch.float16)
    assert sum(int(np.prod(t.shape)) for v in dw.values() for t in v) == sum(int(t.numel()) for v in net.weights.values() for t in v)
    xd = torch.from_numpy(x).cuda()
    inf = net.forward(xd, False)[0]
    scores, tape = net.forward(xd, True)
    ds = torch.empty_like(scores)
    ops.mean_prod(scores, torch.from_numpy(lab).cuda(), torch.zeros(1, device="cuda"), 1.0, 1.0, ds)
    net.ensure_grads().zero_()
    dimg = net.backward(tape, ds, True, True).cpu().numpy()
    torch.cuda.synchronize()
    print("sparse_512: inference rel err %.4f, train rel err %.4f" % (
        _rel(inf.cpu().numpy(), inf_ref), _rel(scores.cpu().numpy(), out.detach().numpy())))
    assert _rel(inf.cpu().numpy(), inf_ref) < 5e-3
    assert _rel(scores.cpu().numpy(), out.detach().numpy()) < 2e-2
    for name in ["discriminator/block_%d/Conv2d" % i for i in range(1, 7)] + ["discriminator/final/Dense_1"]:
        ref = w[name][0].grad.numpy()
        got = net.grads[name][0].cpu().numpy().reshape(ref.shape)
        cos = float(np.vdot(got, ref) / (np.linalg.norm(got) * np.linalg.norm(ref)))
        print("sparse_512 grad %s cos %.4f norm ratio %.4f" % (name, cos, np.linalg.norm(got) / np.linalg.norm(ref)))
        assert cos > 0.97 and abs(np.linalg.norm(got) / np.linalg.norm(ref) - 1) < 0.06
    gi = xt.grad.numpy()
    cos = float(np.vdot(dimg, gi) / (np.linalg.norm(dimg) * np.linalg.norm(gi)))
    print("sparse_512 image gradient cos %.4f norm ratio %.4f" % (cos, np.linalg.norm(dimg) / np.linalg.norm(gi)))
    assert cos > 0.97 and abs(np.linalg.norm(dimg) / np.linalg.norm(gi) - 1) < 0.06


def test_gan2_step_sparse_discriminator_with_output_activation():
    """train_gan3.py:339-354 at BASELINE config 4's shapes (128x128 -> 512x512) with make_discriminator_sparse_512 and
    a tanh output layer (-da tanh) under the relativistic log-sigm loss (-dla log-sigm)."""
    G, D, (gen_train, disc_train, gan_train), ref = _gan_models(True, "relativistic", "log-sigm", disc="sparse", HR=512, dact="tanh")
    x, hr = _data(4, 128, seed=9)
    n = x.shape[0]
    gen = gen_train.predict(x)
    ld = disc_train.train_on_batch([hr, gen], -np.ones(n))
    ld_ref = ref.disc_train_on_batch2(hr, gen)
    print("sparse/tanh disc loss %.5f vs oracle %.5f" % (ld, ld_ref))
    assert abs(ld - ld_ref) <= 3e-2 * abs(ld_ref) + ADV_ATOL
    D.net.set_weights({k: [t.detach().numpy() for t in ts] for k, ts in ref.d.items()})
    lg = gan_train.train_on_batch([x, hr], [hr, -np.ones(n)])
    lg_ref = ref.gan_train_on_batch2(x, hr, hr)
    print("sparse/tanh gan losses %s vs oracle %s" % (np.round(lg, 5), np.round(lg_ref, 5)))
    assert abs(lg[1] - lg_ref[1]) <= 1e-2 * abs(lg_ref[1])
    assert abs(lg[2] - lg_ref[2]) <= 3e-2 * abs(lg_ref[2]) + ADV_ATOL
    # the generator moved the same way: its first conv after the step, against the oracle's
    k = "initial/conv"
    w0 = O.init_upscaler_orig_weights(seed=21, kernel_size=3, res_block_num=1)[k][0]
    ua, ub = G.net.host_weights()[k][0] - w0, ref.g[k][0].detach().numpy() - w0
    agree = float(np.mean(np.sign(ua) == np.sign(ub)))
    print("initial/conv first Adam step: %.3f of the entries moved the same way" % agree)
    assert np.abs(ua).max() < 1.1e-3 and agree > 0.8      # sign-like first step of size lr = 1e-3 (see test_training_gpu)


def test_gan2_disc_loss_fp16_storage():
    """The offset ADV_ATOL allows for is activation-storage rounding, not arithmetic: with fp16 storage (11-bit
    significand) in the critic the offset of the Wasserstein loss of the same batches drops from ~0.025 to 0.005-0.007
    (the generated batch still comes from the bf16 generator)."""
    G, D, (gen_train, disc_train, gan_train), ref = _gan_models(True, "wasserstein", "none", ddtype=torch.float16)
    x, hr = _data(8, HR // 4, seed=6)
    gen = gen_train.predict(x)
    ld = disc_train.train_on_batch([hr, gen], -np.ones(x.shape[0]))
    ld_ref = ref.disc_train_on_batch2(hr, gen)
    print("wasserstein, fp16 critic: disc loss %.5f vs oracle %.5f" % (ld, ld_ref))
    assert abs(ld - ld_ref) <= 1.5e-2
