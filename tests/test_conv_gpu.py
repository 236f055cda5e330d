"""GPU parity: tcgen05 implicit-GEMM conv roles vs the oracle's Keras-semantics layer ops (oracle/keras_ops.py)."""
import numpy as np
import pytest
import torch

from oracle import keras_ops as K
from oracle.dropout_hash import keep_mask

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
    # (N, H, W, Cin, Cout)
    (2, 8, 16, 64, 64),
    (3, 18, 32, 128, 128),
    (2, 36, 64, 64, 256),
    (1, 72, 128, 128, 128),
    (2, 18, 32, 256, 512),
    (1, 4, 8, 64, 128),
]


@pytest.mark.parametrize("n,h,w,cin,cout", CASES)
def test_conv_k3s1_fprop(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _bf16_round(_rand((n, h, w, cin), 1))
    wt = _rand((3, 3, cin, cout), 2, 0.05)
    b = _rand((cout,), 3, 0.1)
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=True)
    y = ops.conv_igemm(ops.CONV_K3S1, x.cuda().to(torch.bfloat16), wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1)
    torch.cuda.synchronize()
    ref = K.leaky_relu(K.conv2d_same(x, _bf16_round(wt), b, 1))
    assert _rel_l2(y.float().cpu(), ref) < 6e-3


@pytest.mark.parametrize("n,h,w,cin,cout", CASES)
def test_conv_k3s1_dgrad(n, h, w, cin, cout):
    """dgrad of Conv2D k3 s1: conv_igemm with flipped/transposed packing == autograd of the oracle conv."""
    from imagegeneration_b200 import ops
    # forward conv maps cout_f=cin -> here: gradient tensor has `cin` channels, produced dx has `cout`
    dy = _bf16_round(_rand((n, h, w, cin), 4))
    wt = _rand((3, 3, cout, cin), 5, 0.05)      # forward weight HWIO with I=cout (dx channels), O=cin (dy channels)
    x = torch.zeros((n, h, w, cout), requires_grad=True)
    K.conv2d_same(x, _bf16_round(wt), None, 1).backward(dy)
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=False, flip=True)
    dx = ops.conv_igemm(ops.CONV_K3S1, dy.cuda().to(torch.bfloat16), wp)
    torch.cuda.synchronize()
    assert _rel_l2(dx.float().cpu(), x.grad) < 6e-3


@pytest.mark.parametrize("n,h,w,cin,cout", CASES)
def test_conv_k4s2_fprop(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _bf16_round(_rand((n, 2 * h, 2 * w, cin), 6))
    wt = _rand((4, 4, cin, cout), 7, 0.05)
    b = _rand((cout,), 8, 0.1)
    wp = ops.pack_conv_weight(ops.CONV_K4S2, wt.cuda(), n_is_d3=True)
    y = ops.conv_igemm(ops.CONV_K4S2, x.cuda().to(torch.bfloat16), wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1)
    torch.cuda.synchronize()
    ref = K.leaky_relu(K.conv2d_same(x, _bf16_round(wt), b, 2))
    assert _rel_l2(y.float().cpu(), ref) < 6e-3


@pytest.mark.parametrize("n,h,w,cin,cout", CASES)
def test_convt_k4s2_fprop(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _bf16_round(_rand((n, h, w, cin), 9))
    wt = _rand((4, 4, cout, cin), 10, 0.05)     # Keras Conv2DTranspose kernel [kh,kw,Cout,Cin]
    beta = _rand((cout,), 11, 0.1)
    wp = ops.pack_conv_weight(ops.CONVT_K4S2, wt.cuda(), n_is_d3=False)
    y = ops.conv_igemm(ops.CONVT_K4S2, x.cuda().to(torch.bfloat16), wp, bias=beta.cuda(), act=ops.ACT_RELU)
    torch.cuda.synchronize()
    ref = torch.relu(K.conv2d_transpose_same(x, _bf16_round(wt), 2) + beta)
    assert _rel_l2(y.float().cpu(), ref) < 6e-3


def test_conv_dropout_matches_hash():
    """The fused dropout uses exactly the mask oracle/dropout_hash.py restates."""
    from imagegeneration_b200 import ops
    n, h, w, cin, cout = 2, 18, 32, 64, 128
    x = _bf16_round(_rand((n, h, w, cin), 12))
    wt = _rand((3, 3, cin, cout), 13, 0.05)
    b = _rand((cout,), 14, 0.1)
    seed = 0x1234_5678_9ABC_DEF1
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=True)
    y = ops.conv_igemm(ops.CONV_K3S1, x.cuda().to(torch.bfloat16), wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1,
                       dropout_thresh8=128, dropout_seed=seed)
    torch.cuda.synchronize()
    pre = K.leaky_relu(K.conv2d_same(x, _bf16_round(wt), b, 1))
    mask = torch.from_numpy(keep_mask(pre.numel(), seed, 128).astype(np.float32)).reshape(pre.shape)
    ref = K.dropout(pre, mask, 0.5)
    yc = y.float().cpu()
    assert torch.equal(yc == 0, (mask == 0) | (ref == 0))
    assert _rel_l2(yc, ref) < 6e-3
    assert abs(float(mask.mean()) - 0.5) < 0.01


def test_conv_dact_epilogue():
    """Backward epilogue: dgrad * LeakyReLU'/dropout mask recovered from the saved forward output."""
    from imagegeneration_b200 import ops
    n, h, w, c = 2, 18, 32, 128
    dy = _bf16_round(_rand((n, h, w, c), 15))
    wt = _rand((3, 3, c, c), 16, 0.05)
    yprev = _bf16_round(_rand((n, h, w, c), 17))
    yprev[torch.rand(yprev.shape, generator=torch.Generator().manual_seed(18)) < 0.5] = 0.0
    x = torch.zeros((n, h, w, c), requires_grad=True)
    K.conv2d_same(x, _bf16_round(wt), None, 1).backward(dy)
    slope = torch.where(yprev > 0, torch.tensor(2.0), torch.where(yprev < 0, torch.tensor(0.2), torch.tensor(0.0)))
    ref = x.grad * slope
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=False, flip=True)
    dx = ops.conv_igemm(ops.CONV_K3S1, dy.cuda().to(torch.bfloat16), wp, yprev=yprev.cuda().to(torch.bfloat16),
                        dact=ops.ACT_LEAKY, dact_alpha=0.1, dact_scale=2.0, dact_zero=0.0)
    torch.cuda.synchronize()
    assert _rel_l2(dx.float().cpu(), ref) < 6e-3
