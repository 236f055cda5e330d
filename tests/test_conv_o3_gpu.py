"""GPU parity: the 64 -> 3 channel gather kernel (csrc/b2g_conv_o3.cu: one 64 -> 27 GEMM per pixel + 9-neighbour
gather over a ring of rows) vs the oracle's Keras-semantics ops and vs the tap kernels it replaces (B2G_CONV_O3=0).
Roles: G.convT4 forward + tanh (sndcgan/SNDCGAN.py:57-61) and D.conv1 input gradient with tanh' (:73-78, :248)."""
import os

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


def _o3_and_taps(fn):
    os.environ["B2G_CONV_O3"] = "1"
    a = fn()
    os.environ["B2G_CONV_O3"] = "0"
    try:
        b = fn()
    finally:
        os.environ.pop("B2G_CONV_O3", None)
    torch.cuda.synchronize()
    return a, b


# rows per CTA of 2 (image boundaries inside and between CTAs), odd heights, one image, both widths
SHAPES = [(2, 6, 256), (1, 9, 128), (3, 5, 256), (5, 2, 128), (7, 1, 256), (1, 4, 128), (2, 144, 256), (40, 9, 128)]


@pytest.mark.parametrize("n,h,w", SHAPES)
def test_o3_forward_tanh_matches_oracle_and_tap_kernel(n, h, w):
    from imagegeneration_b200 import ops
    x = _bf(_rand((n, h, w, 64), 4))
    wt = _rand((3, 3, 3, 64), 5, 0.1)                     # Conv2DTranspose kernel [kh,kw,Cout=3,Cin=64]
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=False, flip=True)
    xg = x.cuda().to(torch.bfloat16)
    y, y_tap = _o3_and_taps(lambda: ops.conv_igemm(ops.CONV_K3S1, xg, wp, act=ops.ACT_TANH, out_f32=True))
    ref = torch.tanh(K.conv2d_transpose_same(x, _bf(wt), 1))
    assert tuple(y.shape) == (n, h, w, 3) and y.dtype == torch.float32
    assert _rel_l2(y, ref) < 2e-3
    assert _rel_l2(y, y_tap) < 1e-4
    # per-row check: a wrong ring slot or image boundary shows up as one bad row, not as a small global error
    err = (y.cpu() - ref).abs().amax(dim=(2, 3))
    assert float(err.max()) < 2e-3, f"worst row {int(err.argmax())}"


@pytest.mark.parametrize("n,h,w", SHAPES[:5])
def test_o3_input_gradient_with_tanh_prime_and_bias(n, h, w):
    from imagegeneration_b200 import ops
    dy = _bf(_rand((n, h, w, 64), 6))
    wc = _rand((3, 3, 3, 64), 7, 0.1)                     # Conv2D kernel [kh,kw,Cin=3,Cout=64]
    xin = torch.zeros((n, h, w, 3), requires_grad=True)
    K.conv2d_same(xin, _bf(wc), None, 1).backward(dy)
    gout = torch.tanh(_rand((n, h, w, 3), 8, 2.0))
    wpd = ops.pack_conv_weight(ops.CONV_K3S1, wc.cuda(), n_is_d3=False, flip=True)
    dyg = dy.cuda().to(torch.bfloat16)
    dx, dx_tap = _o3_and_taps(lambda: ops.conv_igemm(ops.CONV_K3S1, dyg, wpd, yprev=gout.cuda(), dact=ops.ACT_TANH,
                                                     out_f32=True))
    assert _rel_l2(dx, xin.grad * (1 - gout * gout)) < 2e-3
    assert _rel_l2(dx, dx_tap) < 1e-4
    # bias, no activation
    b = _rand((3,), 9, 0.5)
    yb, yb_tap = _o3_and_taps(lambda: ops.conv_igemm(ops.CONV_K3S1, dyg, wpd, bias=b.cuda(), out_f32=True))
    assert _rel_l2(yb, xin.grad + b) < 2e-3
    assert _rel_l2(yb, yb_tap) < 1e-4


def test_o3_full_size_matches_tap_kernel():
    """BASELINE batch 64 at 144 x 256: every CTA walks ~62 rows across image boundaries."""
    from imagegeneration_b200 import ops
    n, h, w = 64, 144, 256
    x = _rand((n, h, w, 64), 11).cuda().to(torch.bfloat16)
    wt = _rand((3, 3, 3, 64), 12, 0.1)
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=False, flip=True)
    y, y_tap = _o3_and_taps(lambda: ops.conv_igemm(ops.CONV_K3S1, x, wp, act=ops.ACT_TANH, out_f32=True))
    assert _rel_l2(y, y_tap) < 1e-4
    assert float((y - y_tap).abs().max()) < 1e-3
