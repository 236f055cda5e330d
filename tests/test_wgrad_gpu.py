"""GPU parity: tcgen05 MN-major weight-gradient kernel vs autograd of the oracle's Keras-semantics layers."""
import pytest
import torch

from oracle import keras_ops as K

pytestmark = pytest.mark.gpu


def _rel_l2(a, b):
    a = a.double().flatten()
    b = b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _bf16_round(t):
    return t.to(torch.bfloat16).to(torch.float32)


def _rand(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(shape, generator=g) * 2 - 1) * scale


CASES = [
    # (N, H(out grid), W(out grid), Cin, Cout)
    (2, 8, 16, 64, 128),
    (3, 18, 32, 128, 128),
    (2, 36, 64, 128, 256),
    (1, 72, 128, 128, 64),
    (2, 18, 32, 256, 512),
    (5, 18, 32, 512, 512),
]


@pytest.mark.parametrize("n,h,w,cin,cout", CASES)
def test_wgrad_conv_k3s1(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _bf16_round(_rand((n, h, w, cin), 1))
    dy = _bf16_round(_rand((n, h, w, cout), 2))
    wt = torch.zeros((3, 3, cin, cout), requires_grad=True)
    K.conv2d_same(x, wt, None, 1).backward(dy)
    dw = torch.zeros((3, 3, cin, cout), device="cuda")
    ops.conv_wgrad(ops.CONV_K3S1, x.cuda().to(torch.bfloat16), dy.cuda().to(torch.bfloat16), dw)
    torch.cuda.synchronize()
    assert _rel_l2(dw.cpu(), wt.grad) < 2e-3


@pytest.mark.parametrize("n,h,w,cin,cout", CASES)
def test_wgrad_conv_k4s2(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _bf16_round(_rand((n, 2 * h, 2 * w, cin), 3))
    dy = _bf16_round(_rand((n, h, w, cout), 4))
    wt = torch.zeros((4, 4, cin, cout), requires_grad=True)
    K.conv2d_same(x, wt, None, 2).backward(dy)
    dw = torch.zeros((4, 4, cin, cout), device="cuda")
    ops.conv_wgrad(ops.CONV_K4S2, x.cuda().to(torch.bfloat16), dy.cuda().to(torch.bfloat16), dw)
    torch.cuda.synchronize()
    assert _rel_l2(dw.cpu(), wt.grad) < 2e-3


@pytest.mark.parametrize("n,h,w,cin,cout", CASES)
def test_wgrad_convt_k4s2(n, h, w, cin, cout):
    """Conv2DTranspose kernel [4,4,Cout,Cin]: shifted operand = dY (Cout channels), unshifted = layer input."""
    from imagegeneration_b200 import ops
    x = _bf16_round(_rand((n, h, w, cin), 5))
    dy = _bf16_round(_rand((n, 2 * h, 2 * w, cout), 6))
    wt = torch.zeros((4, 4, cout, cin), requires_grad=True)
    K.conv2d_transpose_same(x, wt, 2).backward(dy)
    dw = torch.zeros((4, 4, cout, cin), device="cuda")
    ops.conv_wgrad(ops.CONV_K4S2, dy.cuda().to(torch.bfloat16), x.cuda().to(torch.bfloat16), dw)
    torch.cuda.synchronize()
    assert _rel_l2(dw.cpu(), wt.grad) < 2e-3


SLAB_CASES = [
    # (N, H(grid), W(grid) >= 64, Cin, Cout = 128): shapes taken by the slab kernel (b2g_wgrad_slab.cu)
    (2, 9, 64, 128, 128),
    (1, 72, 128, 128, 128),
    (2, 5, 128, 64, 128),
    (3, 7, 64, 256, 128),
]


@pytest.mark.parametrize("n,h,w,cin,cout", SLAB_CASES)
def test_wgrad_slab_k3s1(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _bf16_round(_rand((n, h, w, cin), 11))
    dy = _bf16_round(_rand((n, h, w, cout), 12))
    wt = torch.zeros((3, 3, cin, cout), requires_grad=True)
    K.conv2d_same(x, wt, None, 1).backward(dy)
    dw = torch.zeros((3, 3, cin, cout), device="cuda")
    ops.conv_wgrad(ops.CONV_K3S1, x.cuda().to(torch.bfloat16), dy.cuda().to(torch.bfloat16), dw)
    torch.cuda.synchronize()
    assert _rel_l2(dw.cpu(), wt.grad) < 2e-3
    ops.conv_wgrad(ops.CONV_K3S1, x.cuda().to(torch.bfloat16), dy.cuda().to(torch.bfloat16), dw)   # accumulates
    torch.cuda.synchronize()
    assert _rel_l2(dw.cpu(), 2 * wt.grad) < 2e-3


@pytest.mark.parametrize("n,h,w,cin,cout", SLAB_CASES)
def test_wgrad_slab_k4s2(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _bf16_round(_rand((n, 2 * h, 2 * w, cin), 13))
    dy = _bf16_round(_rand((n, h, w, cout), 14))
    wt = torch.zeros((4, 4, cin, cout), requires_grad=True)
    K.conv2d_same(x, wt, None, 2).backward(dy)
    dw = torch.zeros((4, 4, cin, cout), device="cuda")
    ops.conv_wgrad(ops.CONV_K4S2, x.cuda().to(torch.bfloat16), dy.cuda().to(torch.bfloat16), dw)
    torch.cuda.synchronize()
    assert _rel_l2(dw.cpu(), wt.grad) < 2e-3
