"""GPU parity of the WGAN path (reference wasserstein_gan/WGAN.py) vs the CPU oracle
(oracle/wgan_oracle.py): the new kernels one by one (training-mode BatchNorm forward / backward, RMSprop + clip,
Wasserstein head, LeakyReLU stem) at <= 6e-3 rel-L2 (bf16 storage) or 1e-5 (fp32), then the three train_on_batch
flavours end to end with shared weights and inputs.

End-to-end tolerances: losses within 2e-2 of the logit scale; gradients <= 0.2 rel-L2 against the oracle run with
precision emulation (bf16 activations / weights) — gradients of a LeakyReLU network are not Lipschitz in bf16
perturbations; updated variables are compared through the kernels' own unit tests."""
import ctypes as C

import pytest
import torch

from oracle import keras_ops as K
from oracle import wgan_oracle as W

pytestmark = pytest.mark.gpu


def _rel_l2(a, b):
    a = a.double().flatten().cpu()
    b = b.double().flatten().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _s():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


@pytest.mark.parametrize("P,Cn", [(4 * 16 * 32, 64), (1000, 128), (4 * 2 * 4, 512), (37, 256), (200000, 64)])
def test_batchnorm_training_forward_backward(P, Cn):
    from imagegeneration_b200 import lib as L
    lib = L.load()
    g = torch.Generator().manual_seed(P + Cn)
    x = _bf(torch.randn(P, Cn, generator=g) * 1.5 + 0.3 * torch.randn(Cn, generator=g))
    gamma = torch.rand(Cn, generator=g) + 0.5
    beta = torch.randn(Cn, generator=g) * 0.2
    dy = _bf(torch.randn(P, Cn, generator=g) * 0.01)
    alpha = 0.2
    # oracle (fp64 on the bf16-rounded inputs)
    xd = x.double().requires_grad_(True)
    gd, bd = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    moving = [torch.zeros(Cn, dtype=torch.float64), torch.ones(Cn, dtype=torch.float64)]
    pre = W.batchnorm(xd.reshape(1, 1, P, Cn), gd, bd, moving, True).reshape(P, Cn)
    y = K.leaky_relu(pre, alpha)
    # the CUDA backward takes g = dL/d(pre) (the producing conv epilogue applies LeakyReLU')
    gpre = dy.double()
    dx_ref, dg_ref, db_ref = torch.autograd.grad(pre, (xd, gd, bd), gpre)
    # CUDA
    xc, yc = x.to(torch.bfloat16).cuda(), torch.empty(P, Cn, dtype=torch.bfloat16, device="cuda")
    sums, stat = torch.empty(16 * Cn, device="cuda"), torch.empty(2 * Cn, device="cuda")
    mm, mv = torch.zeros(Cn, device="cuda"), torch.ones(Cn, device="cuda")
    gc, bc = gamma.cuda(), beta.cuda()
    L.check(lib.b2g_bn_stats(_p(xc), _p(sums), P, Cn, _s()), "stats")
    L.check(lib.b2g_bn_finalize(_p(sums), P, Cn, 1e-3, 0.99, _p(stat), _p(mm), _p(mv), None, _s()), "fin")
    L.check(lib.b2g_bn_apply_act(_p(xc), _p(stat), _p(gc), _p(bc), alpha, _p(yc), P, Cn, _s()), "apply")
    torch.cuda.synchronize()
    mean_ref = x.double().mean(0)
    var_ref = x.double().var(0, unbiased=False)
    assert _rel_l2(stat[:Cn], mean_ref) < 1e-4 or float((stat[:Cn].cpu().double() - mean_ref).abs().max()) < 1e-5
    assert _rel_l2(stat[Cn:], 1.0 / torch.sqrt(var_ref + 1e-3)) < 1e-4
    assert _rel_l2(mm, moving[0]) < 1e-4 or float((mm.cpu().double() - moving[0]).abs().max()) < 1e-6
    assert _rel_l2(mv, moving[1]) < 1e-5
    assert _rel_l2(yc.float(), y) < 6e-3
    dyc = dy.to(torch.bfloat16).cuda()
    dxc = torch.empty_like(dyc)
    red, dgc, dbc = torch.empty(16 * Cn, device="cuda"), torch.empty(Cn, device="cuda"), torch.empty(Cn, device="cuda")
    L.check(lib.b2g_bn_bwd(_p(dyc), _p(xc), _p(stat), _p(gc), _p(red), _p(dxc), _p(dgc), _p(dbc), P, Cn, _s()), "bwd", 2)
    torch.cuda.synchronize()
    assert _rel_l2(dgc, dg_ref) < 1e-3
    assert _rel_l2(dbc, db_ref) < 1e-3
    assert _rel_l2(dxc.float(), dx_ref) < 8e-3
    # in place (dx over g) gives the same result up to the summation order of the atomics in the reduction pass
    L.check(lib.b2g_bn_bwd(_p(dyc), _p(xc), _p(stat), _p(gc), _p(red), _p(dyc), None, None, P, Cn, _s()), "bwd", 2)
    torch.cuda.synchronize()
    assert _rel_l2(dyc.float(), dxc.float()) < 2e-3
    # inference statistics
    L.check(lib.b2g_bn_inference_stat(_p(mm), _p(mv), Cn, 1e-3, _p(stat), None, _s()), "inf")
    L.check(lib.b2g_bn_apply_act(_p(xc), _p(stat), _p(gc), _p(bc), alpha, _p(yc), P, Cn, _s()), "apply")
    torch.cuda.synchronize()
    yi = K.leaky_relu(W.batchnorm(x.double(), gamma.double(), beta.double(), moving, False), alpha)
    assert _rel_l2(yc.float(), yi) < 6e-3


@pytest.mark.parametrize("kind,shape,cout", [("k3s1", (2, 16, 32, 64), 128), ("k4s2", (2, 16, 32, 64), 128),
                                             ("convt", (2, 8, 16, 128), 64), ("k3s1", (1, 8, 128, 64), 64),
                                             ("convt", (1, 4, 64, 128), 64), ("k3s1", (2, 8, 16, 256), 256),
                                             ("k4s2", (2, 8, 128, 64), 128), ("c3", (2, 16, 32, 3), 64),
                                             ("c3", (1, 8, 256, 3), 64)])
def test_conv_forward_accumulates_batch_statistics(kind, shape, cout):
    """colsum + colsumsq on a forward call (no bias): the per-channel sum and sum of squares of the STORED bf16
    outputs, from every conv kernel family (per-tap, CTA pair, row slab, 3-channel edge), TMA and direct epilogues."""
    from imagegeneration_b200 import ops
    g = torch.Generator().manual_seed(sum(shape) + cout)
    n, h, w, cin = shape
    if kind == "c3":
        x = (torch.rand(shape, generator=g) * 2 - 1).cuda()
        wt = (torch.randn((3, 3, 3, cout), generator=g) * 0.2).cuda()
        wp = ops.pack_conv_weight(ops.CONV_K1S1, wt.view(1, 1, 27, cout), n_is_d3=True, ck_pad=64)
        k = ops.CONV_K3S1_C3
    else:
        x = (torch.rand(shape, generator=g) * 2 - 1).to(torch.bfloat16).cuda()
        if kind == "convt":
            wt = (torch.randn((4, 4, cout, cin), generator=g) * 0.05).cuda()
            wp, k = ops.pack_conv_weight(ops.CONVT_K4S2, wt, n_is_d3=False), ops.CONVT_K4S2
        elif kind == "k4s2":
            wt = (torch.randn((4, 4, cin, cout), generator=g) * 0.05).cuda()
            wp, k = ops.pack_conv_weight(ops.CONV_K4S2, wt, n_is_d3=True), ops.CONV_K4S2
        else:
            wt = (torch.randn((3, 3, cin, cout), generator=g) * 0.05).cuda()
            wp, k = ops.pack_conv_weight(ops.CONV_K3S1, wt, n_is_d3=True), ops.CONV_K3S1
    ref = ops.conv_igemm(k, x, wp, act=ops.ACT_NONE)
    cs, cq = torch.zeros(cout, device="cuda"), torch.zeros(cout, device="cuda")
    out = ops.conv_igemm(k, x, wp, act=ops.ACT_NONE, colsum=cs, colsumsq=cq)
    torch.cuda.synchronize()
    assert torch.equal(out, ref)
    o = out.double().reshape(-1, cout)
    assert _rel_l2(cs, o.sum(0)) < 1e-4 or float((cs.double().cpu() - o.sum(0).cpu()).abs().max()) < 1e-3
    assert _rel_l2(cq, (o * o).sum(0)) < 1e-5


def test_batchnorm_rejects_bad_shapes():
    from imagegeneration_b200 import lib as L
    lib = L.load()
    x = torch.zeros(8, 24, dtype=torch.bfloat16, device="cuda")
    s = torch.zeros(16 * 64, device="cuda")
    assert lib.b2g_bn_stats(_p(x), _p(s), 8, 24, _s()) != 0        # 24 channels: not a power of two
    assert lib.b2g_bn_stats(None, _p(s), 8, 64, _s()) != 0
    assert b"b2g_bn_stats" in lib.b2g_last_error()


def test_rmsprop_with_clip():
    from imagegeneration_b200 import lib as L
    lib = L.load()
    g = torch.Generator().manual_seed(11)
    n, clip_n = 4096 + 8, 1024
    w = torch.randn(n, generator=g) * 0.02
    gr = torch.randn(n, generator=g) * 1e-3
    rms = torch.rand(n, generator=g) * 1e-6
    wc, gc, rc = w.cuda(), (gr * 2).cuda(), rms.cuda()
    wo, ro = w.clone().double(), rms.clone().double()
    for _ in range(3):
        L.check(lib.b2g_rmsprop(_p(wc), _p(gc), _p(rc), n, 5e-5, 0.9, 1e-7, 0.5, clip_n, 0.01, _s()), "rmsprop")
        W.rmsprop_step(wo, gr.double(), ro, 5e-5)
        wo[:clip_n].clamp_(-0.01, 0.01)
    torch.cuda.synchronize()
    assert _rel_l2(wc, wo) < 1e-6 and _rel_l2(rc, ro) < 1e-6
    assert float(wc[:clip_n].abs().max()) <= 0.01 and float(wc[clip_n:].abs().max()) > 0.01
    assert lib.b2g_rmsprop(_p(wc), _p(gc), _p(rc), n - 1, 5e-5, 0.9, 1e-7, 1.0, 0, 0.0, _s()) != 0


IMAGE = (16, 32, 3)
BATCH = 4


def _make_pair(emu):
    from imagegeneration_b200.wgan import WGANEngine
    eng = WGANEngine(IMAGE, BATCH, 2, seed=62)
    orc = W.OracleWGAN(IMAGE, seed=62, emu=emu)
    g = torch.Generator().manual_seed(5)
    cv = [v.detach().cpu().clone() for v in eng.critic_model.trainable_variables]
    gv = [v.detach().cpu().clone() for v in eng.generator_model.trainable_variables]
    for lst in (cv, gv):                      # non-trivial biases / gamma / beta
        for v in lst:
            if v.dim() == 1:
                v.add_((torch.rand(v.shape, generator=g) - 0.5) * 0.2)
    eng.critic_model.set_weights(cv)
    eng.generator_model.set_weights(gv)
    orc.cv, orc.gv = [v.clone() for v in cv], [v.clone() for v in gv]
    orc.c_rms = [torch.zeros_like(v) for v in orc.cv]
    orc.gan_rms_g = [torch.zeros_like(v) for v in orc.gv]
    return eng, orc


def _inputs(seed=0):
    g = torch.Generator().manual_seed(77 + seed)
    real = torch.rand((BATCH, *IMAGE), generator=g) * 2 - 1
    z = torch.randn((BATCH, 128), generator=g)
    return real, z


def test_variable_order_matches_keras():
    from imagegeneration_b200.wgan import WGANEngine
    eng = WGANEngine(IMAGE, BATCH, 5)
    assert [tuple(v.shape) for v in eng.critic_model.trainable_variables] == [s for _, s in W.critic_var_shapes(IMAGE)]
    assert [tuple(v.shape) for v in eng.generator_model.trainable_variables] == [s for _, s in W.generator_var_shapes(IMAGE)]
    c = eng.critic_model
    assert c.clip_n == sum(v.numel() for v in c.arena.variables[:7])      # conv kernels are multiples of 4 elements


def test_critic_train_on_batch_matches_oracle():
    eng, orc = _make_pair(emu=True)
    real, _z = _inputs()
    col = {}
    cv_before = [v.clone() for v in orc.cv]
    lo = orc.train_critic_on_batch(real, -1.0, col)
    le = eng.train_critic_on_batch(real.cuda(), -1.0)
    torch.cuda.synchronize()
    scale = float(col["logits"].abs().mean()) + 1e-6
    assert abs(float(le) - float(lo)) < 2e-2 * scale, (float(le), float(lo))
    C_ = eng.critic_model
    for i in range(7):
        # the engine's raw conv output leaves the bias out (BatchNorm cancels it; bn_finalize adds it to the moving mean)
        raw_with_bias = C_._raws[i].float() + C_.arena.variables[7 + i]
        assert _rel_l2(raw_with_bias, col[f"raw{i + 1}"]) < 2e-2, f"raw{i + 1}"
        assert _rel_l2(C_._acts[i].float(), col[f"c{i + 1}"]) < 2e-2, f"c{i + 1}"
        assert _rel_l2(C_.bn.moving_var[i], orc.c_moving[i][1]) < 1e-3
        assert float((C_.bn.moving_mean[i].cpu() - orc.c_moving[i][0]).abs().max()) < 1e-3
    names = [n for n, _ in W.critic_var_shapes(IMAGE)]
    for n, ge, go in zip(names, C_.trainable_gradients, col["grads"]):
        if n.startswith("conv") and n.endswith("bias"):
            continue            # cancelled by BatchNorm: rounding noise on both sides
        assert _rel_l2(ge, go) < 0.2, n
    # the update itself, checked exactly on the engine's own gradients: RMSprop from zero slots, then the clip
    for n, ve, ge, vb in zip(names, C_.trainable_variables, C_.trainable_gradients, cv_before):
        gd = ge.cpu().double()
        want = vb.double() - 5e-5 * gd / (torch.sqrt(0.1 * gd * gd) + 1e-7)
        if n.endswith("kernel") and n.startswith("conv"):
            want = want.clamp(-W.CLIP, W.CLIP)
            assert float(ve.abs().max()) <= W.CLIP + 1e-9
        assert float((ve.cpu().double() - want).abs().max()) < 1e-6, n
    assert eng.critic_optimizer.iterations == 1


def test_gan_train_on_batch_matches_oracle():
    eng, orc = _make_pair(emu=True)
    _real, z = _inputs(1)
    col = {}
    lo = orc.train_gan_on_batch(z, col)
    cvars_before = [v.clone() for v in eng.critic_model.arena.variables[:16]]
    le = eng.train_gan_on_batch(z.cuda())
    torch.cuda.synchronize()
    assert _rel_l2(eng.generator_model.g_out, col["fake"]) < 2e-2
    logit_scale = float(eng.critic_model.logits.abs().mean()) + 1e-6
    # the loss is the mean of 4 logits of magnitude ~0.5 and mixed sign (|loss| ~ 0.04): 4e-2 of the logit scale
    # (measured 1e-2 ... 2.1e-2, varying with the summation order of the reductions' atomics)
    assert abs(float(le) - float(lo)) < 4e-2 * logit_scale, (float(le), float(lo))
    gnames = [n for n, _ in W.generator_var_shapes(IMAGE)]
    # 0.35: at this size the oracle's OWN fp32 and bf16-emulating runs differ by 0.23-0.29 rel-L2 in these gradients
    # (LeakyReLU mask flips + BatchNorm over 32-2048 values per channel); measured here 0.10-0.26
    for n, ge, go in zip(gnames, eng.generator_model.arena.grads, col["g_grads"]):
        assert _rel_l2(ge, go) < 0.35, n
    cg = eng.critic_model.trainable_gradients
    for j, i in enumerate(orc.critic_bn_indices()):
        assert _rel_l2(cg[i], col["cbn_grads"][j]) < 0.35, f"critic bn grad {i}"
    # frozen critic conv / Dense variables did not move; BatchNorm gamma/beta did
    for a, b in zip(eng.critic_model.arena.variables[:16], cvars_before):
        assert torch.equal(a, b)
    assert eng.gan_model.optimizer.iterations == 1 and eng.critic_optimizer.iterations == 0
    assert float(eng.gan_model.rms_cbn.abs().sum()) > 0


def test_predict_uses_moving_statistics():
    eng, orc = _make_pair(emu=False)
    _real, z = _inputs(2)
    # move the moving statistics away from (0, 1) on both sides
    g = torch.Generator().manual_seed(9)
    for i, c in enumerate((256, 128, 64)):
        m, v = torch.randn(c, generator=g) * 0.01, torch.rand(c, generator=g) * 0.5 + 0.01
        orc.g_moving[i] = [m.clone(), v.clone()]
        eng.generator_model.bn.moving_mean[i].copy_(m)
        eng.generator_model.bn.moving_var[i].copy_(v)
    a = eng.generate_fake_samples(z.cuda())
    b = orc.predict(z)
    assert _rel_l2(a, b) < 2e-2


def _sync(eng, orc):
    """teacher forcing: copy the oracle's variables and moving statistics into the engine"""
    eng.critic_model.set_weights(orc.cv)
    eng.generator_model.set_weights(orc.gv)
    for bn, mv in ((eng.critic_model.bn, orc.c_moving), (eng.generator_model.bn, orc.g_moving)):
        for i, (m, v) in enumerate(mv):
            bn.moving_mean[i].copy_(m)
            bn.moving_var[i].copy_(v)


@pytest.mark.parametrize("emu", [True, False])
def test_train_iteration_teacher_forced_losses(emu):
    """Every sub-step of two WGAN.train iterations (WGAN.py:296-318) starting from the ORACLE's evolving state
    (variables + moving statistics copied in before each sub-step).

    emu=True  (oracle with the CUDA path's storage precision: bf16 activations / packed weights): every loss within
              2e-2 of the logit scale (measured <= 7e-3) — the engine computes exactly the restated arithmetic.
    emu=False (plain fp32 oracle): within 0.25 of the logit scale (measured 4e-3 ... 0.14).  At this size the problem
              is ill-conditioned for ANY reduced precision: the untrained generator's images differ by 0.004 (std over
              the batch) around a 0.016 pattern, the deep BatchNorms see 32 values per channel (4 images x 2 x 4) and
              renormalise those differences to unit variance, amplifying bf16 rounding; recorded in DESIGN.md."""
    eng, orc = _make_pair(emu=emu)
    tol = 2e-2 if emu else 0.25
    g = torch.Generator().manual_seed(123)
    reals = [torch.rand((BATCH, *IMAGE), generator=g) * 2 - 1 for _ in range(2)]
    zs = [torch.randn((BATCH, 128), generator=g) for _ in range(3)]

    def close(a, b, what):
        scale = float(eng.critic_model.logits.abs().mean()) + 1e-3
        assert abs(float(a) - float(b)) < tol * scale, (what, float(a), float(b), scale)

    for it in range(2):
        for x, z in zip(reals, zs[:2]):
            _sync(eng, orc)
            close(eng.train_critic_on_batch(x.cuda(), -1.0), orc.train_critic_on_batch(x, -1.0), (it, "real"))
            _sync(eng, orc)
            fe, fo = eng.generate_fake_samples(z.cuda()), orc.predict(z)
            assert _rel_l2(fe, fo) < 2e-2
            close(eng.train_critic_on_batch(fo.cuda(), 1.0), orc.train_critic_on_batch(fo, 1.0), (it, "fake"))
        _sync(eng, orc)
        close(eng.train_gan_on_batch(zs[2].cuda()), orc.train_gan_on_batch(zs[2]), (it, "gan"))


def test_train_iteration_free_running():
    """Free-running iterations: step counts, the clip invariant, finite losses, and losses that stay near the
    precision-emulating oracle.  RMSprop's first steps are sign-like (lr*g/sqrt(0.1 g^2)): weights whose gradient is
    at bf16-noise level step in either direction, so trajectories separate much faster than with Adam/SGD — the bound
    here is deliberately loose (0.35 of the loss scale); the tight checks are the teacher-forced test above and the
    exact-update checks of the single-step tests."""
    eng, orc = _make_pair(emu=True)
    g = torch.Generator().manual_seed(123)
    reals = [torch.rand((BATCH, *IMAGE), generator=g) * 2 - 1 for _ in range(2)]
    zs = [torch.randn((BATCH, 128), generator=g) for _ in range(3)]
    for it in range(2):
        ro = orc.train_iteration(reals, zs[:2], zs[2])
        re = eng.train_iteration([r.cuda() for r in reals], [z.cuda() for z in zs[:2]], zs[2].cuda())
        torch.cuda.synchronize()
        scale = max(abs(float(x)) for x in ro) + 1e-3
        for name, a, b in zip(("c1", "c2", "g"), re, ro):
            assert torch.isfinite(a).all()
            assert abs(float(a) - float(b)) < 0.35 * scale, (it, name, float(a), float(b))
    assert eng.critic_optimizer.iterations == 8 and eng.gan_model.optimizer.iterations == 2
    for i in range(7):
        assert float(eng.critic_model.arena.variables[i].abs().max()) <= W.CLIP + 1e-9


def test_wgan_class_trains_over_a_dataset():
    from imagegeneration_b200.wgan import WGAN
    g = torch.Generator().manual_seed(4)
    data = [torch.rand((BATCH, *IMAGE), generator=g) * 2 - 1 for _ in range(4)]
    wg = WGAN(data, IMAGE, BATCH, 2)
    hist = wg.train(1)
    assert len(hist["g_hist"]) == 2 and len(hist["c1_hist"]) == 2 and wg.epoch == 1
    assert all(torch.isfinite(torch.tensor(v)).all() for v in hist.values())


def test_ragged_last_batch_and_empty_batch():
    """image_dataset_from_directory does not drop the remainder: the last real batch of an epoch can be smaller than
    bach_size while generate_fake_samples still draws bach_size fakes (WGAN.py:298-303).  An empty batch is an error."""
    from imagegeneration_b200.lib import B200GanError
    from imagegeneration_b200.wgan import WGAN
    g = torch.Generator().manual_seed(8)
    data = [torch.rand((b, *IMAGE), generator=g) * 2 - 1 for b in (BATCH, BATCH, BATCH, 3, 1)]
    wg = WGAN(data, IMAGE, BATCH, 5)
    hist = wg.train(1)
    assert len(hist["g_hist"]) == 1 and wg.critic_optimizer.iterations == 10, (hist, wg.critic_optimizer.iterations)
    assert all(torch.isfinite(torch.tensor(v)).all() for v in hist.values()), hist
    # same ragged batch through the oracle: the batch-of-3 critic step still matches (3 images: the deep BatchNorms
    # normalise 24 values per channel, so the bound is wider than the 2e-2 of the batch-4 tests; measured <= 1e-2)
    eng, orc = _make_pair(emu=True)
    x3 = data[3]
    le, lo = eng.train_critic_on_batch(x3.cuda(), -1.0), orc.train_critic_on_batch(x3, -1.0)
    scale = float(eng.critic_model.logits.abs().mean()) + 1e-3
    assert abs(float(le) - float(lo)) < 1e-1 * scale, (float(le), float(lo), scale)
    with pytest.raises(B200GanError):
        eng.train_critic_on_batch(torch.zeros((0, *IMAGE), device="cuda"), -1.0)


def test_cpu_tensors_are_rejected():
    from imagegeneration_b200.lib import B200GanError
    from imagegeneration_b200.wgan import WGANEngine
    eng = WGANEngine(IMAGE, BATCH, 5)
    with pytest.raises(B200GanError):
        eng.train_critic_on_batch(torch.zeros((BATCH, *IMAGE)), -1.0)
