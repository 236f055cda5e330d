"""CPU: the oracle restatement itself — Keras-2 layer semantics (documented behaviour; the reference has no tests,
so parity with TensorFlow is UNPINNED), algorithmic FLOP counts of SURVEY.md §8, the frozen fixtures under
tests/golden/, and the bit-exact dropout hash shared with the CUDA kernels."""
import importlib.util
import math
import os

import numpy as np
import torch
import torch.nn.functional as F

from oracle import keras_ops as K
from oracle import sndcgan_oracle as O
from oracle.dropout_hash import dropout_bits, keep_mask, mix32, rate_to_thresh8

HERE = os.path.dirname(os.path.abspath(__file__))


def test_golden_fixtures_reproduce():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    got = mod.run()
    ref = np.load(os.path.join(HERE, "golden", "oracle_tiny.npz"))
    for k in ref.files:
        if k.startswith("hash"):
            assert np.array_equal(got[k], ref[k]), k
        else:
            assert np.allclose(got[k], ref[k], rtol=1e-6, atol=1e-9), k


def test_dropout_hash_known_answers():
    # lowbias32 reference values (the public test vectors of the hash: 0 -> 0, 1 -> 0x6a6a3ec4 ... computed once)
    assert int(mix32(np.array([0], dtype=np.uint32))[0]) == 0
    bits = dropout_bits(np.arange(4, dtype=np.uint32), 0)
    assert bits.dtype == np.uint32 and len(set(bits.tolist())) == 4
    m = keep_mask(1 << 16, 12345, 128)
    assert abs(m.mean() - 0.5) < 0.01
    m2 = keep_mask(1 << 16, 12346, 128)
    assert abs((m == m2).mean() - 0.5) < 0.02             # different seeds decorrelate
    assert keep_mask(64, 1, 0).all()
    assert rate_to_thresh8(0.5) == 128 and rate_to_thresh8(0.0) == 0 and rate_to_thresh8(1.0) == 255


def test_conv_semantics():
    g = torch.Generator().manual_seed(0)
    x = torch.randn((2, 6, 8, 3), generator=g, dtype=torch.float64)
    w = torch.randn((3, 3, 3, 5), generator=g, dtype=torch.float64)
    # SAME, stride 1: output pixel (y,x) = sum_k in(y+ky-1, x+kx-1) w[ky,kx]  (direct loop)
    y = K.conv2d_same(x, w, None, 1)
    ref = torch.zeros_like(y)
    xp = F.pad(x, (0, 0, 1, 1, 1, 1))
    for ky in range(3):
        for kx in range(3):
            ref += torch.einsum("nhwc,cd->nhwd", xp[:, ky:ky + 6, kx:kx + 8, :], w[ky, kx])
    assert torch.allclose(y, ref)
    # SAME, k4 s2 on even sizes: pad 1 both sides
    w4 = torch.randn((4, 4, 3, 5), generator=g, dtype=torch.float64)
    y2 = K.conv2d_same(x, w4, None, 2)
    assert tuple(y2.shape) == (2, 3, 4, 5)
    ref2 = torch.zeros_like(y2)
    for ky in range(4):
        for kx in range(4):
            ref2 += torch.einsum("nhwc,cd->nhwd", xp[:, ky:ky + 6:2, kx:kx + 8:2, :], w4[ky, kx])
    assert torch.allclose(y2, ref2)
    # Conv2DTranspose == gradient of Conv2D w.r.t. its input (TF's definition), kernel [kh,kw,Cout,Cin]
    xin = torch.zeros((2, 6, 8, 3), dtype=torch.float64, requires_grad=True)
    dy = torch.randn((2, 3, 4, 5), generator=g, dtype=torch.float64)
    K.conv2d_same(xin, w4, None, 2).backward(dy)
    assert torch.allclose(K.conv2d_transpose_same(dy, w4, 2), xin.grad)


def test_loss_bn_adam_semantics():
    x = torch.tensor([[-3.0], [0.0], [2.5]], dtype=torch.float64)
    for label in (0.0, 1.0):
        ref = F.binary_cross_entropy_with_logits(x, torch.full_like(x, label))
        assert abs(float(K.bce_with_logits(x, label)) - float(ref)) < 1e-12
    gamma, beta = torch.tensor([2.0]), torch.tensor([0.5])
    assert abs(float(K.batchnorm_inference(torch.tensor([3.0]), gamma, beta)) - (2 * 3 / math.sqrt(1.001) + 0.5)) < 1e-6
    # Keras Adam, two steps on a scalar, by hand (epsilon outside the bias-corrected sqrt)
    w, m, v = torch.tensor([1.0], dtype=torch.float64), torch.zeros(1, dtype=torch.float64), torch.zeros(1, dtype=torch.float64)
    g = torch.tensor([0.1], dtype=torch.float64)
    K.adam_step(w, g, m, v, 1, 2e-4)
    lr1 = 2e-4 * math.sqrt(1 - 0.999) / (1 - 0.9)
    # Keras forms (1 - beta) in the variable dtype, fp32: 1 - float32(0.9), 1 - float32(0.999)
    c1, c2 = float(np.float32(1) - np.float32(0.9)), float(np.float32(1) - np.float32(0.999))
    assert abs(c2 - 0.001) / 0.001 > 1e-5                                      # the fp32 constant is visibly not 0.001
    assert abs(float(w) - (1.0 - lr1 * 0.1 * c1 / (math.sqrt(0.01 * c2) + 1e-7))) < 1e-12
    K.adam_step(w, g, m, v, 2, 2e-4)
    m2 = 0.1 * c1 + (0.1 - 0.1 * c1) * c1
    v2 = 0.01 * c2 + (0.01 - 0.01 * c2) * c2
    assert abs(float(m) - m2) < 1e-12 and abs(float(v) - v2) < 1e-12
    assert abs(float(K.leaky_relu(torch.tensor([-2.0]))) + 0.2) < 1e-7
    assert float(K.dropout(torch.ones(4), torch.tensor([1.0, 0, 1, 0]), 0.5).sum()) == 4.0


def test_variable_inventory_and_flops_match_the_survey():
    gs = O.generator_var_shapes()
    ds = O.discriminator_var_shapes()
    assert len(gs) == 13 and len(ds) == 16                                    # SURVEY §2.1 (Adam row)
    assert sum(int(np.prod(s)) for _, s in gs) == 41_093_696
    assert sum(int(np.prod(s)) for _, s in ds) == 6_147_585
    f = O.flops_per_image()
    assert f["g_fwd"] == 2 * 3_725_328_384 and f["d_fwd"] == 2 * 7_764_738_048
    gen = torch.Generator().manual_seed(1)
    w = K.glorot_uniform((3, 3, 128, 128), gen)
    assert float(w.abs().max()) <= math.sqrt(6 / (9 * 128 * 2)) and float(w.abs().max()) > 0.9 * math.sqrt(6 / (9 * 256))


def test_step_ordering_and_adam_iterations():
    """one train_step = 1 G update + 2 D updates; the fake batch of D step 2 is the stale g_output."""
    orc = O.OracleSNDCGAN((16, 32, 3), dropout=0.0, seed=3)
    real = torch.rand(2, 16, 32, 3) * 2 - 1
    z = torch.rand(2, 128) * 2 - 1
    g_before = [v.clone() for v in orc.gv]
    out = orc.train_step(real, z)
    assert orc.g_opt.iterations == 1 and orc.d_opt.iterations == 2
    assert abs(float(out[1]) - float(out[2]) - float(out[3])) < 1e-6           # d_loss = real + fake (:264)
    assert tuple(out[4].shape) == (2, 1) and tuple(out[5].shape) == (2, 1)
    # the logits returned for the fake batch come from G_t (pre-update) weights
    with torch.no_grad():
        stale = O.generator_forward(g_before, z, (16, 32, 3))
        fresh = O.generator_forward(orc.gv, z, (16, 32, 3))
        l_stale = O.discriminator_forward(orc.dv, stale, 0.0, None)
    assert not torch.allclose(stale, fresh)
    # D has been updated once more after producing out[5], so only check the shape/finite-ness relation here
    assert torch.isfinite(l_stale).all()


def test_spectral_norm_oracle_vs_svd():
    g = torch.Generator().manual_seed(5)
    w = torch.randn((300, 40), generator=g, dtype=torch.float64)
    w += 30 * torch.outer(torch.randn(300, generator=g, dtype=torch.float64), torch.randn(40, generator=g, dtype=torch.float64)) / 100
    u = torch.randn(40, generator=g, dtype=torch.float64)
    u /= u.norm()
    for _ in range(200):
        sigma, u, v = K.spectral_norm_power_iteration(w, u, 1)
    assert abs(float(sigma) - float(torch.linalg.svdvals(w)[0])) / float(sigma) < 1e-8
