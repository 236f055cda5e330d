"""GPU parity: the row-slab conv kernel (csrc/b2g_conv_rows.cu; layers >= 128 pixels wide, Cout <= 128) vs the oracle's
Keras-semantics layer ops, and vs the per-tap kernel it replaces on those layers (B2G_CONV_ROWS=0)."""
import os

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


def _handled(kind, n, h, w, cin, bn):
    import ctypes as C
    from imagegeneration_b200 import lib
    out = (C.c_int * 464)()
    assert lib.load().b2g_conv_rows_plan_query(kind, n, h, w, cin, bn, 0, out, 464) == 0
    return out[0] == 1


def _both(fn):
    """Runs fn() with the row-slab kernel and with the per-tap kernel."""
    os.environ["B2G_CONV_ROWS"] = "1"
    a = fn()
    os.environ["B2G_CONV_ROWS"] = "0"
    try:
        b = fn()
    finally:
        os.environ["B2G_CONV_ROWS"] = "1"
    torch.cuda.synchronize()
    return a, b


WIDE = [
    # (N, H, W, Cin, Cout): W >= 128, odd heights exercise the masked last row block
    (2, 5, 128, 64, 64),
    (1, 8, 256, 64, 128),
    (2, 7, 128, 128, 64),
    (1, 72, 128, 128, 128),
    (3, 2, 128, 192, 128),
    (2, 1, 128, 64, 64),
]


@pytest.mark.parametrize("n,h,w,cin,cout", WIDE)
def test_rows_k3s1_fprop_and_dgrad(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    assert _handled(ops.CONV_K3S1, n, h, w, cin, 128 if cout % 128 == 0 else 64)
    x = _bf16_round(_rand((n, h, w, cin), 1))
    wt = _rand((3, 3, cin, cout), 2, 0.05)
    b = _rand((cout,), 3, 0.1)
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=True)
    xg = x.cuda().to(torch.bfloat16)
    y, y_tap = _both(lambda: ops.conv_igemm(ops.CONV_K3S1, xg, wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1))
    ref = K.leaky_relu(K.conv2d_same(x, _bf16_round(wt), b, 1))
    assert _rel_l2(y.float().cpu(), ref) < 6e-3
    assert _rel_l2(y.float(), y_tap.float()) < 2e-3
    # dgrad role: flipped / transposed packing
    dy = _bf16_round(_rand((n, h, w, cout), 4))
    xt = torch.zeros((n, h, w, cin), requires_grad=True)
    K.conv2d_same(xt, _bf16_round(wt), None, 1).backward(dy)
    wpd = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=False, flip=True)
    dx = ops.conv_igemm(ops.CONV_K3S1, dy.cuda().to(torch.bfloat16), wpd)
    torch.cuda.synchronize()
    assert _rel_l2(dx.float().cpu(), xt.grad) < 6e-3


@pytest.mark.parametrize("n,h,w,cin,cout", WIDE)
def test_rows_k4s2_fprop(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    assert _handled(ops.CONV_K4S2, n, 2 * h, 2 * w, cin, 128 if cout % 128 == 0 else 64)
    x = _bf16_round(_rand((n, 2 * h, 2 * w, cin), 6))
    wt = _rand((4, 4, cin, cout), 7, 0.05)
    b = _rand((cout,), 8, 0.1)
    wp = ops.pack_conv_weight(ops.CONV_K4S2, wt.cuda(), n_is_d3=True)
    xg = x.cuda().to(torch.bfloat16)
    y, y_tap = _both(lambda: ops.conv_igemm(ops.CONV_K4S2, xg, wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1))
    ref = K.leaky_relu(K.conv2d_same(x, _bf16_round(wt), b, 2))
    assert _rel_l2(y.float().cpu(), ref) < 6e-3
    assert _rel_l2(y.float(), y_tap.float()) < 2e-3


@pytest.mark.parametrize("n,h,w,cin,cout", WIDE)
def test_rows_convt_fprop(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    assert _handled(ops.CONVT_K4S2, n, h, w, cin, 128 if cout % 128 == 0 else 64)
    x = _bf16_round(_rand((n, h, w, cin), 9))
    wt = _rand((4, 4, cout, cin), 10, 0.05)
    beta = _rand((cout,), 11, 0.1)
    wp = ops.pack_conv_weight(ops.CONVT_K4S2, wt.cuda(), n_is_d3=False)
    xg = x.cuda().to(torch.bfloat16)
    y, y_tap = _both(lambda: ops.conv_igemm(ops.CONVT_K4S2, xg, wp, bias=beta.cuda(), act=ops.ACT_RELU))
    ref = torch.relu(K.conv2d_transpose_same(x, _bf16_round(wt), 2) + beta)
    assert _rel_l2(y.float().cpu(), ref) < 6e-3
    assert _rel_l2(y.float(), y_tap.float()) < 2e-3


def test_rows_dropout_and_dact_epilogues():
    from imagegeneration_b200 import ops
    n, h, w, cin, cout = 2, 6, 128, 64, 128
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
    # backward epilogue on the convT role (D.conv2 input gradient): LeakyReLU'/dropout from the saved output
    dy = _bf16_round(_rand((n, h, w, 128), 15))
    wt2 = _rand((4, 4, 64, 128), 16, 0.05)      # Conv2D k4s2 HWIO 64 -> 128; its dgrad is a convT 128 -> 64
    yprev = _bf16_round(_rand((n, 2 * h, 2 * w, 64), 17))
    yprev[torch.rand(yprev.shape, generator=torch.Generator().manual_seed(18)) < 0.5] = 0.0
    xt = torch.zeros((n, 2 * h, 2 * w, 64), requires_grad=True)
    K.conv2d_same(xt, _bf16_round(wt2), None, 2).backward(dy)
    slope = torch.where(yprev > 0, torch.tensor(2.0), torch.where(yprev < 0, torch.tensor(0.2), torch.tensor(0.0)))
    wpd = ops.pack_conv_weight(ops.CONVT_K4S2, wt2.cuda(), n_is_d3=False)
    dx = ops.conv_igemm(ops.CONVT_K4S2, dy.cuda().to(torch.bfloat16), wpd, yprev=yprev.cuda().to(torch.bfloat16),
                        dact=ops.ACT_LEAKY, dact_alpha=0.1, dact_scale=2.0, dact_zero=0.0)
    torch.cuda.synchronize()
    assert _rel_l2(dx.float().cpu(), xt.grad * slope) < 6e-3


@pytest.mark.parametrize("n,h,w", [(2, 6, 256), (1, 9, 128)])
def test_rows_narrow_fp32_output_with_tanh(n, h, w):
    """G.convT4 forward (64 -> 3 channels, tanh): k3s1 with the 32-wide N tile, 4 rows per super-tile."""
    from imagegeneration_b200 import ops
    assert _handled(ops.CONV_K3S1, n, h, w, 64, 32)
    x = _bf16_round(_rand((n, h, w, 64), 19))
    wt = _rand((3, 3, 3, 64), 20, 0.05)           # Conv2DTranspose k3 s1 kernel [kh,kw,Cout=3,Cin=64]
    ref = torch.tanh(K.conv2d_transpose_same(x, _bf16_round(wt), 1))
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=False, flip=True)
    xg = x.cuda().to(torch.bfloat16)
    os.environ["B2G_CONV_O3"] = "0"              # this test is about the tap kernels the gather kernel replaces
    try:
        y, y_tap = _both(lambda: ops.conv_igemm(ops.CONV_K3S1, xg, wp, act=ops.ACT_TANH, out_f32=True))
    finally:
        os.environ.pop("B2G_CONV_O3", None)
    assert tuple(y.shape) == (n, h, w, 3)
    assert _rel_l2(y.cpu(), ref) < 2e-3
    assert _rel_l2(y, y_tap) < 1e-4


# ---------------------------------------------------------------------------------------------------------------
# CTA-pair kernel (csrc/b2g_conv_pair.cu, tcgen05 cta_group::2): same results as the one-CTA kernels
# ---------------------------------------------------------------------------------------------------------------
def _pair_vs_single(fn):
    os.environ["B2G_CONV_PAIR"] = "1"
    a = fn()
    os.environ["B2G_CONV_PAIR"] = "0"
    try:
        b = fn()
    finally:
        os.environ["B2G_CONV_PAIR"] = "1"
    torch.cuda.synchronize()
    return a, b


PAIR = [
    # (kind, N, H, W, Cin, Cout): even number of 128-pixel tiles; Cout 64 / 128 / 256 / 512 (two N tiles)
    ("k3s1", 2, 8, 16, 64, 64),
    ("k3s1", 4, 18, 32, 128, 256),
    ("k3s1", 2, 18, 32, 256, 512),
    ("k4s2", 2, 16, 32, 64, 128),
    ("k4s2", 2, 36, 64, 128, 256),
    ("convt", 2, 8, 16, 128, 64),
    ("convt", 4, 18, 32, 256, 256),
    ("k3s1", 3, 36, 64, 64, 128),      # 3*36*64/128 = 54 tiles
]


@pytest.mark.parametrize("kind,n,h,w,cin,cout", PAIR)
def test_pair_kernel_matches_single_cta_and_oracle(kind, n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    b = _rand((cout,), 33, 0.1)
    seed = 0x77_1234_5678
    if kind == "k3s1":
        x = _bf16_round(_rand((n, h, w, cin), 31))
        wt = _rand((3, 3, cin, cout), 32, 0.05)
        wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=True)
        k, ref = ops.CONV_K3S1, K.leaky_relu(K.conv2d_same(x, _bf16_round(wt), b, 1))
    elif kind == "k4s2":
        x = _bf16_round(_rand((n, 2 * h, 2 * w, cin), 31))
        wt = _rand((4, 4, cin, cout), 32, 0.05)
        wp = ops.pack_conv_weight(ops.CONV_K4S2, wt.cuda(), n_is_d3=True)
        k, ref = ops.CONV_K4S2, K.leaky_relu(K.conv2d_same(x, _bf16_round(wt), b, 2))
    else:
        x = _bf16_round(_rand((n, h, w, cin), 31))
        wt = _rand((4, 4, cout, cin), 32, 0.05)
        wp = ops.pack_conv_weight(ops.CONVT_K4S2, wt.cuda(), n_is_d3=False)
        k, ref = ops.CONVT_K4S2, K.leaky_relu(K.conv2d_transpose_same(x, _bf16_round(wt), 2) + b)
    xg = x.cuda().to(torch.bfloat16)
    os.environ["B2G_CONV_ROWS"] = "0"
    try:
        y, y1 = _pair_vs_single(lambda: ops.conv_igemm(k, xg, wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1,
                                                       dropout_thresh8=64, dropout_seed=seed))
    finally:
        os.environ["B2G_CONV_ROWS"] = "1"
    assert torch.equal(y, y1)            # same taps in the same order into fp32 accumulators: bit-identical
    mask = torch.from_numpy(keep_mask(ref.numel(), seed, 64).astype(np.float32)).reshape(ref.shape)
    assert _rel_l2(y.float().cpu(), K.dropout(ref, mask, 0.25)) < 6e-3
