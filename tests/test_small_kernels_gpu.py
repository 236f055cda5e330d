"""GPU parity of the edge-layer, dense, head, optimizer and spectral-norm kernels vs oracle/keras_ops.py."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

from oracle import keras_ops as K

pytestmark = pytest.mark.gpu


def _rel_l2(a, b):
    a = a.double().flatten().cpu()
    b = b.double().flatten().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _bf(t):
    return t.to(torch.bfloat16).to(torch.float32)


def _rand(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(shape, generator=g) * 2 - 1) * scale


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _s():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


EDGE = [(2, 16, 32), (1, 144, 256), (3, 24, 64)]


@pytest.mark.parametrize("n,h,w", EDGE)
def test_edge_c3_to_c64_tensor_core(n, h, w):
    """D.conv1 forward (SNDCGAN.py:73-78) = im2col27 + K1S1 GEMM; input is rounded to bf16 by the im2col."""
    from imagegeneration_b200 import ops
    x = _rand((n, h, w, 3), 1)
    wt = _rand((3, 3, 3, 64), 2, 0.2)
    b = _rand((64,), 3, 0.1)
    x27 = ops.im2col27(x.cuda())
    wp = ops.pack_conv_weight(ops.CONV_K1S1, wt.cuda().view(1, 1, 27, 64), n_is_d3=True, ck_pad=64)
    y = ops.conv_igemm(ops.CONV_K1S1, x27, wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1)
    torch.cuda.synchronize()
    ref = K.leaky_relu(K.conv2d_same(_bf(x), _bf(wt), b, 1))
    assert _rel_l2(y.float(), ref) < 6e-3
    # the ones channel and the zero padding of the im2col
    assert float(x27[..., 27].float().min()) == 1.0 and float(x27[..., 28:].float().abs().max()) == 0.0


@pytest.mark.parametrize("n,h,w", EDGE)
def test_edge_c64_to_c3_tensor_core(n, h, w):
    """G.convT4 forward + tanh (SNDCGAN.py:57-61): K3S1 with flipped taps, narrow fp32 output."""
    from imagegeneration_b200 import ops
    x = _bf(_rand((n, h, w, 64), 4))
    wt = _rand((3, 3, 3, 64), 5, 0.1)                     # Conv2DTranspose kernel [kh,kw,Cout=3,Cin=64]
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=False, flip=True)
    y = ops.conv_igemm(ops.CONV_K3S1, x.cuda().to(torch.bfloat16), wp, act=ops.ACT_TANH, out_f32=True)
    torch.cuda.synchronize()
    ref = torch.tanh(K.conv2d_transpose_same(x, _bf(wt), 1))
    assert tuple(y.shape) == (n, h, w, 3) and y.dtype == torch.float32
    assert _rel_l2(y, ref) < 2e-3
    # D.conv1 input-gradient with tanh' of the generator output
    dy = _bf(_rand((n, h, w, 64), 6))
    wc = _rand((3, 3, 3, 64), 7, 0.1)                     # Conv2D kernel [kh,kw,Cin=3,Cout=64]
    xin = torch.zeros((n, h, w, 3), requires_grad=True)
    K.conv2d_same(xin, _bf(wc), None, 1).backward(dy)
    gout = torch.tanh(_rand((n, h, w, 3), 8, 2.0))
    wpd = ops.pack_conv_weight(ops.CONV_K3S1, wc.cuda(), n_is_d3=False, flip=True)
    dx = ops.conv_igemm(ops.CONV_K3S1, dy.cuda().to(torch.bfloat16), wpd, yprev=gout.cuda(), dact=ops.ACT_TANH,
                        out_f32=True)
    torch.cuda.synchronize()
    assert _rel_l2(dx, xin.grad * (1 - gout * gout)) < 2e-3


@pytest.mark.parametrize("n,h,w", EDGE)
def test_edge_wgrad_tensor_core(n, h, w):
    """[3,3,3,64] weight gradient + bias gradient through the im2col27 operand (ones channel = bias row)."""
    from imagegeneration_b200 import ops
    x = _rand((n, h, w, 3), 9)
    dy = _bf(_rand((n, h, w, 64), 10))
    wt = torch.zeros((3, 3, 3, 64), requires_grad=True)
    bb = torch.zeros((64,), requires_grad=True)
    K.conv2d_same(_bf(x), wt, bb, 1).backward(dy)
    x27 = ops.im2col27(x.cuda())
    dw = torch.zeros((3, 3, 3, 64), device="cuda")
    db = torch.zeros((64,), device="cuda")
    ops.conv_wgrad(ops.CONV_K1S1, x27, dy.cuda().to(torch.bfloat16), dw, ca_valid=27, dbias=db)
    torch.cuda.synchronize()
    assert _rel_l2(dw, wt.grad) < 2e-3
    assert _rel_l2(db, bb.grad) < 2e-3


@pytest.mark.parametrize("n,h,w", EDGE)
def test_edge_wgrad_fused_patch_kernel(n, h, w):
    """Same gradients with the patch built inside the kernel (kind K3S1_C3): no im2col27 tensor; bit-compatible
    accumulation semantics (dw / dbias are added to)."""
    from imagegeneration_b200 import ops
    x = _rand((n, h, w, 3), 9)
    dy = _bf(_rand((n, h, w, 64), 10))
    wt = torch.zeros((3, 3, 3, 64), requires_grad=True)
    bb = torch.zeros((64,), requires_grad=True)
    K.conv2d_same(_bf(x), wt, bb, 1).backward(dy)
    dw = torch.zeros((3, 3, 3, 64), device="cuda")
    db = torch.zeros((64,), device="cuda")
    ops.conv_wgrad(ops.CONV_K3S1_C3, x.cuda(), dy.cuda().to(torch.bfloat16), dw, dbias=db)
    torch.cuda.synchronize()
    assert _rel_l2(dw, wt.grad) < 2e-3
    assert _rel_l2(db, bb.grad) < 2e-3
    ops.conv_wgrad(ops.CONV_K3S1_C3, x.cuda(), dy.cuda().to(torch.bfloat16), dw)      # accumulates, bias optional
    torch.cuda.synchronize()
    assert _rel_l2(dw, 2 * wt.grad) < 2e-3 and _rel_l2(db, bb.grad) < 2e-3


@pytest.mark.parametrize("n,h,w", EDGE[:2])
def test_edge_cuda_core_variants(n, h, w):
    """The direct (CUDA-core) edge kernels kept in the ABI agree with the oracle too."""
    from imagegeneration_b200 import lib as L
    lib = L.load()
    x = _rand((n, h, w, 3), 11)
    wt = _rand((3, 3, 3, 64), 12, 0.2)
    b = _rand((64,), 13, 0.1)
    out = torch.empty((n, h, w, 64), dtype=torch.bfloat16, device="cuda")
    xc, wc, bc = x.cuda(), wt.cuda(), b.cuda()
    L.check(lib.b2g_conv3x3_c3_to_c64(_p(xc), _p(wc), None, _p(bc), _p(out), n, h, w, L.ACT_LEAKY, 0.1, 0, 0, None,
                                      L.ACT_NONE, _s()), "c3_to_c64")
    torch.cuda.synchronize()
    assert _rel_l2(out.float(), K.leaky_relu(K.conv2d_same(x, wt, b, 1))) < 6e-3
    x64 = _bf(_rand((n, h, w, 64), 14))
    o3 = torch.empty((n, h, w, 3), device="cuda")
    x64c = x64.cuda().to(torch.bfloat16)
    L.check(lib.b2g_conv3x3_c64_to_c3(_p(x64c), _p(wc), None, _p(o3), n, h, w, L.ACT_TANH, None, L.ACT_NONE, _s()),
            "c64_to_c3")
    torch.cuda.synchronize()
    assert _rel_l2(o3, torch.tanh(K.conv2d_transpose_same(x64, wt, 1))) < 1e-4
    dy = _bf(_rand((n, h, w, 64), 15))
    w0 = torch.zeros((3, 3, 3, 64), requires_grad=True)
    b0 = torch.zeros((64,), requires_grad=True)
    K.conv2d_same(x, w0, b0, 1).backward(dy)
    dw = torch.zeros((3, 3, 3, 64), device="cuda")
    db = torch.zeros((64,), device="cuda")
    dyc = dy.cuda().to(torch.bfloat16)
    L.check(lib.b2g_wgrad3x3_c3_c64(_p(xc), _p(dyc), _p(dw), _p(db), n, h, w, _s()), "wgrad3x3")
    torch.cuda.synchronize()
    assert _rel_l2(dw, w0.grad) < 1e-4 and _rel_l2(db, b0.grad) < 1e-4


@pytest.mark.parametrize("B,F", [(4, 4096), (64, 512 * 6 * 8), (33, 2048)])
def test_dense_generator_stem(B, F):
    """Dense(128->F, no bias) + inference BatchNorm + ReLU and its backward (SNDCGAN.py:31-33)."""
    from imagegeneration_b200 import lib as L
    lib = L.load()
    inv_std = 1.0 / math.sqrt(1.0 + 1e-3)
    z = _rand((B, 128), 20)
    W = _rand((128, F), 21, 0.05).requires_grad_(True)
    gamma = (1 + _rand((F,), 22, 0.2)).requires_grad_(True)
    beta = _rand((F,), 23, 0.1).requires_grad_(True)
    ref = torch.relu(K.batchnorm_inference(K.dense(z, W), gamma, beta))
    out = torch.empty((B, F), dtype=torch.bfloat16, device="cuda")
    zc, Wc, gc, bc = z.cuda(), W.detach().cuda(), gamma.detach().cuda(), beta.detach().cuda()
    L.check(lib.b2g_dense_g_fwd(_p(zc), _p(Wc), _p(gc), _p(bc), inv_std, _p(out), B, F, _s()), "dense fwd")
    torch.cuda.synchronize()
    assert _rel_l2(out.float(), ref) < 4e-3
    dy = _bf(_rand((B, F), 24)) * (ref > 0)              # gradient w.r.t. the BN output, ReLU' applied
    ref.backward(dy)
    dW = torch.empty((128, F), device="cuda")
    dg = torch.empty((F,), device="cuda")
    db = torch.empty((F,), device="cuda")
    dq = dy.cuda().to(torch.bfloat16)
    L.check(lib.b2g_dense_g_bwd(_p(zc), _p(Wc), _p(gc), inv_std, _p(dq), _p(dW), _p(dg), _p(db), B, F, _s()), "dense bwd")
    torch.cuda.synchronize()
    assert _rel_l2(dW, W.grad) < 1e-4 and _rel_l2(dg, gamma.grad) < 1e-4 and _rel_l2(db, beta.grad) < 1e-4


@pytest.mark.parametrize("B,F", [(4, 4096), (64, 294912)])
@pytest.mark.parametrize("kind,label", [(0, 1.0), (0, 0.0), (1, 1.0), (1, 0.0), (2, 1.0)])
def test_dhead_loss_and_backward(B, F, kind, label):
    """Flatten + Dense(1) + BCE-with-logits (SNDCGAN.py:122-123,170) or hinge, and the head backward."""
    from imagegeneration_b200 import lib as L
    lib = L.load()
    x = _bf(_rand((B, F), 30))
    x[torch.rand(x.shape, generator=torch.Generator().manual_seed(31)) < 0.5] = 0.0
    w = _rand((F, 1), 32, 0.02).requires_grad_(True)
    b = torch.tensor([0.05], requires_grad=True)
    logits = K.dense(x, w, b)
    if kind == 0:
        loss = K.bce_with_logits(logits, label)
    elif kind == 1:
        loss = K.hinge_d_loss(logits, label)
    else:
        loss = -logits.mean()
    loss.backward()
    xc, wc, bc = x.cuda().to(torch.bfloat16), w.detach().cuda(), b.detach().cuda()
    ws = torch.empty(64 * B, device="cuda")
    lg = torch.empty((B, 1), device="cuda")
    dl = torch.empty(B, device="cuda")
    lo = torch.empty(2, device="cuda")
    L.check(lib.b2g_dhead_fwd_loss(_p(xc), _p(wc), None, _p(bc), B, F, kind, label, 1.0 / B, _p(ws), ws.numel(), _p(lg),
                                   _p(dl), _p(lo), _s()), "head fwd", 2)
    dx = torch.empty((B, F), dtype=torch.bfloat16, device="cuda")
    dw = torch.zeros((F, 1), device="cuda")
    db = torch.zeros(1, device="cuda")
    L.check(lib.b2g_dhead_bwd(_p(xc), _p(wc), None, _p(dl), B, F, 0.1, 2.0, 0.0, _p(dx), _p(dw), _p(db), _s()), "head bwd")
    torch.cuda.synchronize()
    assert float((lg.cpu() - logits.detach()).abs().max()) < 1e-3
    assert abs(float(lo[0]) - float(loss)) < 1e-4 * (1 + abs(float(loss)))
    assert abs(float(lo[1]) - float(torch.sigmoid(logits).mean())) < 1e-4
    assert _rel_l2(dw, w.grad) < 1e-4 and abs(float(db) - float(b.grad)) < 1e-5
    slope = torch.where(x > 0, torch.tensor(2.0), torch.where(x < 0, torch.tensor(0.2), torch.tensor(0.0)))
    ref_dx = dl.cpu()[:, None] * w.detach().t() * slope
    assert _rel_l2(dx.float(), ref_dx) < 4e-3


def test_adam_matches_keras_formula():
    from imagegeneration_b200 import lib as L
    lib = L.load()
    n = 1 << 20
    w = _rand((n,), 40)
    g = _rand((n,), 41, 1e-3)
    m = torch.zeros(n)
    v = torch.zeros(n)
    wc, mc, vc = w.cuda(), m.cuda(), v.cuda()
    for t in range(1, 4):
        gt = g * t
        K.adam_step(w, gt, m, v, t, 2e-4)
        lr_t = 2e-4 * math.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        gc = (gt * 2).cuda()   # gscale 0.5 undoes the factor 2 (the data-parallel 1/world path)
        L.check(lib.b2g_adam(_p(wc), _p(gc), _p(mc), _p(vc), n, lr_t, 0.9, 0.999, 1e-7, 0.5, _s()), "adam")
    torch.cuda.synchronize()
    assert float((wc.cpu() - w).abs().max()) < 1e-6
    assert _rel_l2(mc, m) < 1e-4 and _rel_l2(vc, v) < 1e-4


@pytest.mark.parametrize("P,Cc", [(1000, 64), (36864 * 2, 128), (777, 512), (4096, 256)])
def test_colsum(P, Cc):
    from imagegeneration_b200 import lib as L
    lib = L.load()
    x = _bf(_rand((P, Cc), 50))
    out = torch.zeros(Cc, device="cuda")
    xc = x.cuda().to(torch.bfloat16)
    L.check(lib.b2g_colsum_bf16(_p(xc), _p(out), P, Cc, _s()), "colsum")
    torch.cuda.synchronize()
    assert float((out.cpu() - x.sum(0)).abs().max()) < 1e-3 * (1 + float(x.sum(0).abs().max()))


def test_bn_fold_finalize():
    from imagegeneration_b200 import lib as L
    lib = L.load()
    taps, Co, Ci = 16, 64, 128
    inv_std = 1.0 / math.sqrt(1.001)
    W = _rand((4, 4, Co, Ci), 60, 0.1)
    G = _rand((4, 4, Co, Ci), 61, 0.1)
    gamma = 1 + _rand((Co,), 62, 0.3)
    Gc, dg = G.cuda(), torch.empty(Co, device="cuda")
    Wc, gc = W.cuda(), gamma.cuda()
    L.check(lib.b2g_bn_fold_finalize(_p(Wc), _p(Gc), _p(gc), inv_std, _p(dg), taps, Co, Ci, _s()), "bn fold")
    torch.cuda.synchronize()
    assert _rel_l2(dg, inv_std * (W * G).sum(dim=(0, 1, 3))) < 1e-5
    assert _rel_l2(Gc, G * (gamma * inv_std)[None, None, :, None]) < 1e-6


def test_spectral_norm_power_iteration_and_grad():
    """sigma after one power iteration within 1e-5 of the oracle (same u0), converged sigma vs fp64 SVD, and the
    gradient through W/sigma (Miyato et al. 2018, Alg. 1; not in the reference)."""
    from imagegeneration_b200 import lib as L
    lib = L.load()
    shapes = [(27, 64), (1024, 128), (1152, 128), (2048, 256), (2304, 256), (4096, 512), (4608, 512), (294912, 1)]
    Ws = [_rand(s, 70 + i, 0.05) for i, s in enumerate(shapes)]
    g = torch.Generator().manual_seed(99)
    for i, s in enumerate(shapes):   # plant a dominant singular direction so the power iteration has a spectral gap
        a, b = torch.randn(s[0], generator=g), torch.randn(s[1], generator=g)
        Ws[i] = Ws[i] + 3.0 * float(torch.linalg.matrix_norm(Ws[i], 2)) * torch.outer(a / a.norm(), b / b.norm())
    us = []
    for s in shapes:
        u = torch.randn(s[1], generator=g)
        us.append(u / u.norm())
    Wc = [w.cuda() for w in Ws]
    uc = [u.clone().cuda() for u in us]
    vr = [torch.zeros(s[0], device="cuda") for s in shapes]
    tr = [torch.zeros(s[1], device="cuda") for s in shapes]
    sc = [torch.zeros(8, device="cuda") for s in shapes]
    arr = lambda ts: (C.c_void_p * 8)(*[t.data_ptr() for t in ts])
    Ks = (C.c_int * 8)(*[s[0] for s in shapes])
    Cs = (C.c_int * 8)(*[s[1] for s in shapes])

    def iterate():
        L.check(lib.b2g_spectral_norm_multi(8, arr(Wc), arr(uc), arr(vr), arr(tr), arr(sc), Ks, Cs, _s()), "sn", 2)

    iterate()
    torch.cuda.synchronize()
    for i in range(8):
        sigma, u1, v1 = K.spectral_norm_power_iteration(Ws[i].double(), us[i].double(), 1)
        assert abs(float(sc[i][1]) - float(sigma)) / float(sigma) < 1e-5, f"layer {i}"
        assert _rel_l2(uc[i], u1) < 1e-4
        assert abs(float(sc[i][2]) * float(sigma) - 1.0) < 1e-5
    # gradient through W/sigma for one layer
    i = 3
    W = Ws[i].clone().requires_grad_(True)
    _sg, u1, v1 = K.spectral_norm_power_iteration(Ws[i], us[i], 1)
    sigma = v1 @ (W @ u1)
    Gsn = _rand(shapes[i], 80, 0.01)
    ((W / sigma) * Gsn).sum().backward()
    Gc = Gsn.cuda()
    dot = torch.zeros(1, device="cuda")
    L.check(lib.b2g_sn_grad_finalize(_p(Wc[i]), _p(Gc), _p(uc[i]), _p(vr[i]), _p(sc[i]), _p(dot), shapes[i][0],
                                     shapes[i][1], _s()), "sn grad", 3)
    torch.cuda.synchronize()
    assert _rel_l2(Gc, W.grad) < 1e-4
    # convergence to the largest singular value
    for _ in range(60):
        iterate()
    torch.cuda.synchronize()
    for i in range(7):
        s0 = float(torch.linalg.svdvals(Ws[i].double())[0])
        assert abs(float(sc[i][1]) - s0) / s0 < 1e-3, f"layer {i}"


@pytest.mark.parametrize("n,h,w", EDGE + [(2, 5, 128), (5, 3, 8)])
def test_edge_c3_fused_patch_kernel(n, h, w):
    """B2G_CONV_K3S1_C3: the im2col patch is built in shared memory; must equal the im2col27 + K1S1 route bit for bit
    (same bf16 operands, same fp32 accumulation order is not guaranteed -> tiny tolerance) and match the oracle."""
    from imagegeneration_b200 import ops
    x = _rand((n, h, w, 3), 21)
    wt = _rand((3, 3, 3, 64), 22, 0.2)
    b = _rand((64,), 23, 0.1)
    wp = ops.pack_conv_weight(ops.CONV_K1S1, wt.cuda().view(1, 1, 27, 64), n_is_d3=True, ck_pad=64)
    seed = 0x0BAD_5EED_1234
    y = ops.conv_igemm(ops.CONV_K3S1_C3, x.cuda(), wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1,
                       dropout_thresh8=128, dropout_seed=seed)
    y2 = ops.conv_igemm(ops.CONV_K1S1, ops.im2col27(x.cuda()), wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1,
                        dropout_thresh8=128, dropout_seed=seed)
    torch.cuda.synchronize()
    assert _rel_l2(y.float(), y2.float()) < 1e-3
    pre = K.leaky_relu(K.conv2d_same(_bf(x), _bf(wt), b, 1))
    from oracle.dropout_hash import keep_mask
    import numpy as np
    mask = torch.from_numpy(keep_mask(pre.numel(), seed, 128).astype(np.float32)).reshape(pre.shape)
    assert _rel_l2(y.float().cpu(), K.dropout(pre, mask, 0.5)) < 6e-3
    # G.convT4 input gradient: 3-channel fp32 dY -> 64 channels, ReLU' of the saved layer input in the epilogue
    dy = _rand((n, h, w, 3), 24)
    wg = _rand((3, 3, 3, 64), 25, 0.2)                     # Conv2DTranspose kernel [kh,kw,Cout=3,Cin=64]
    a3 = torch.relu(_rand((n, h, w, 64), 26))
    xin = a3.clone().requires_grad_(True)
    K.conv2d_transpose_same(xin, _bf(wg), 1).backward(_bf(dy))
    wpd = ops.pack_conv_weight(ops.CONV_K1S1, wg.cuda().view(1, 1, 27, 64), n_is_d3=True, ck_pad=64)
    dq = ops.conv_igemm(ops.CONV_K3S1_C3, dy.cuda(), wpd, yprev=a3.cuda().to(torch.bfloat16), dact=ops.ACT_RELU)
    torch.cuda.synchronize()
    assert _rel_l2(dq.float().cpu(), xin.grad * (a3 > 0).float()) < 6e-3


def test_pack_plan_matches_single_packs():
    """b2g_pack_conv_weight_multi (one launch, all layers) is bit-identical to the per-tensor packing."""
    from imagegeneration_b200 import ops
    g = torch.Generator().manual_seed(5)
    sc = torch.full((1,), 0.37, device="cuda")
    jobs = [
        (ops.CONV_K3S1, (3, 3, 64, 128), dict(n_is_d3=True, scale_scalar=sc)),
        (ops.CONV_K3S1, (3, 3, 64, 128), dict(n_is_d3=False, flip=True)),
        (ops.CONV_K4S2, (4, 4, 128, 64), dict(n_is_d3=True)),
        (ops.CONVT_K4S2, (4, 4, 64, 128), dict(n_is_d3=False, scale_scalar=sc, scale_d2=torch.rand(64, generator=g).cuda())),
        (ops.CONV_K1S1, (1, 1, 27, 64), dict(n_is_d3=True, ck_pad=64)),
        (ops.CONV_K3S1, (3, 3, 3, 64), dict(n_is_d3=False, flip=True, scale_d3=torch.rand(64, generator=g).cuda())),
    ]
    plan = ops.PackPlan()
    ws, outs = [], []
    for kind, shape, kw in jobs:
        w = (torch.rand(shape, generator=g) - 0.5).cuda()
        ws.append(w)
        outs.append(plan.add(kind, w, **kw))
    plan.run()
    torch.cuda.synchronize()
    for (kind, _shape, kw), w, o in zip(jobs, ws, outs):
        ref = ops.pack_conv_weight(kind, w, **kw)
        assert o.shape == ref.shape and torch.equal(o, ref)
    ws[0].mul_(2.0)                 # masters change in place; run() re-packs from them
    plan.run()
    kind, _shape, kw = jobs[0]
    assert torch.equal(outs[0], ops.pack_conv_weight(kind, ws[0], **kw))


def test_adam_fused_with_spectral_norm():
    """b2g_sn_dot_multi + b2g_adam_sn (spectral norm riding on the optimizer pass, north_star item 2) on the real
    discriminator arena layout (8 matrices + biases) vs fp64: gradient through W/sigma, Keras Adam, and the NEXT power
    iteration on the updated weights (sigma <= 1e-5 relative, Miyato et al. Alg. 1: one iteration per update)."""
    from imagegeneration_b200 import lib as L
    lib = L.load()
    shapes = [(27, 64), (1024, 128), (1152, 128), (2048, 256), (2304, 256), (4096, 512), (4608, 512), (2304, 1)]
    offs, pos = [], 0
    for k, c in shapes:                       # kernel, then its bias (cout floats, padded to 4) like ParamArena
        offs.append(pos)
        pos += (k * c + 3) // 4 * 4
        pos += (c + 3) // 4 * 4
    n = pos
    w = _rand((n,), 200, 0.05)
    gen = torch.Generator().manual_seed(201)
    for (k, c), o in zip(shapes, offs):       # spectral gap so that one iteration is well conditioned
        a, b = torch.randn(k, generator=gen), torch.randn(c, generator=gen)
        w[o:o + k * c] += (0.5 * torch.outer(a / a.norm(), b / b.norm())).flatten()
    g = _rand((n,), 202, 0.01)
    m = _rand((n,), 203, 0.003)
    v = _rand((n,), 204, 1.0).abs() * 1e-4
    us = []
    for k, c in shapes:
        u = torch.randn(c, generator=gen)
        us.append(u / u.norm())
    wc, gc, mc, vc = w.cuda(), g.cuda(), m.cuda(), v.cuda()
    uc = [u.clone().cuda() for u in us]
    vr = [torch.zeros(s[0], device="cuda") for s in shapes]
    tr = [torch.zeros(s[1], device="cuda") for s in shapes]
    sc = [torch.zeros(8, device="cuda") for s in shapes]
    dots = torch.zeros(8, device="cuda")
    arr = lambda ts: (C.c_void_p * 8)(*[t.data_ptr() for t in ts])
    Ks = (C.c_int * 8)(*[s[0] for s in shapes])
    Cs = (C.c_int * 8)(*[s[1] for s in shapes])
    off_c = (C.c_long * 8)(*offs)
    Wv = [wc[o:o + k * c] for (k, c), o in zip(shapes, offs)]
    # the iteration the "forward pass" used
    L.check(lib.b2g_spectral_norm_multi(8, arr(Wv), arr(uc), arr(vr), arr(tr), arr(sc), Ks, Cs, _s()), "sn", 2)
    torch.cuda.synchronize()
    u1 = [u.double().cpu() for u in uc]
    v1 = [(x / x.norm()).double().cpu() for x in vr]
    sig1 = [float(s[1]) for s in sc]
    lr_t, b1, b2, eps, gscale = 3e-4, 0.9, 0.999, 1e-7, 0.5
    L.check(lib.b2g_sn_dot_multi(_p(wc), _p(gc), 8, off_c, Ks, Cs, _p(dots), _s()), "dot")
    L.check(lib.b2g_adam_sn(_p(wc), _p(gc), _p(mc), _p(vc), n, 8, off_c, Ks, Cs, arr(uc), arr(vr), arr(tr), arr(sc),
                            _p(dots), lr_t, None, b1, b2, eps, gscale, _s()), "adam_sn")
    torch.cuda.synchronize()
    assert float(dots.abs().max()) == 0.0                     # consumed and re-zeroed
    # reference in fp64
    wd, gd, md, vd = w.double(), g.double().clone(), m.double(), v.double()
    for i, ((k, c), o) in enumerate(zip(shapes, offs)):
        W = wd[o:o + k * c].reshape(k, c)
        G = gd[o:o + k * c].reshape(k, c)
        dot = float((W * G).sum())
        gd[o:o + k * c] = (G / sig1[i] - dot / sig1[i] ** 2 * torch.outer(v1[i], u1[i])).flatten()
    gd *= gscale
    md = md + (gd - md) * (1 - b1)
    vd = vd + (gd * gd - vd) * (1 - b2)
    w_new = wd - lr_t * md / (vd.sqrt() + eps)
    assert _rel_l2(mc, md) < 1e-5 and _rel_l2(vc, vd) < 1e-5
    assert _rel_l2(wc.double().cpu() - wd, w_new - wd) < 2e-4    # the update itself (fp32 rounding of w dominates)
    assert _rel_l2(wc, w_new) < 1e-6
    for i, ((k, c), o) in enumerate(zip(shapes, offs)):
        Wn = w_new[o:o + k * c].reshape(k, c)
        sigma, u2, v2 = K.spectral_norm_power_iteration(Wn, u1[i], 1)
        assert abs(float(sc[i][1]) - float(sigma)) / float(sigma) < 1e-5, f"sigma layer {i}"
        assert _rel_l2(uc[i], u2) < 1e-4 and _rel_l2(vr[i] / vr[i].norm(), v2) < 1e-4, f"u/v layer {i}"
        assert float(tr[i].abs().max()) == 0.0 and float(sc[i][0]) == 0.0   # scratch re-zeroed for the next call
