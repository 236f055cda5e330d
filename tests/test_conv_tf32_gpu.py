"""GPU parity of the fp32 / tf32 path (north_star: per-layer outputs and input gradients within 1e-3 relative L2 of
the fp32 reference): tcgen05 kind::tf32 implicit-GEMM conv roles with fp32 activations, fp32 packed weights and fp32
outputs, compared with the oracle's Keras-semantics layer ops in fp32 on UNROUNDED inputs."""
import numpy as np
import pytest
import torch

from oracle import keras_ops as K
from oracle.dropout_hash import keep_mask

pytestmark = pytest.mark.gpu
TOL = 1e-3      # BASELINE.json north_star: fp32/tf32 path, relative L2 per layer


def _rel_l2(a, b):
    a = a.double().flatten()
    b = b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _rand(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(shape, generator=g) * 2 - 1) * scale


CASES = [
    # (N, H, W, Cin, Cout)
    (2, 8, 16, 32, 64),
    (3, 18, 32, 128, 128),
    (2, 36, 64, 64, 256),
    (1, 72, 128, 128, 128),
    (2, 18, 32, 256, 512),
]


@pytest.mark.parametrize("n,h,w,cin,cout", CASES)
def test_tf32_k3s1_fprop_and_dgrad(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _rand((n, h, w, cin), 1)
    wt = _rand((3, 3, cin, cout), 2, 0.05)
    b = _rand((cout,), 3, 0.1)
    wp = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=True, tf32=True)
    y = ops.conv_igemm(ops.CONV_K3S1, x.cuda(), wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1, tf32=True)
    torch.cuda.synchronize()
    assert y.dtype == torch.float32
    assert _rel_l2(y.cpu(), K.leaky_relu(K.conv2d_same(x, wt, b, 1))) < TOL
    # input gradient (flipped / transposed packing) with the LeakyReLU / dropout derivative of the saved output
    if cin % 64 == 0:      # the produced tensor (dx) needs a multiple of 64 channels
        dy = _rand((n, h, w, cout), 4)
        xt = torch.zeros((n, h, w, cin), requires_grad=True)
        K.conv2d_same(xt, wt, None, 1).backward(dy)
        yprev = _rand((n, h, w, cin), 5)
        yprev[torch.rand(yprev.shape, generator=torch.Generator().manual_seed(6)) < 0.5] = 0.0
        slope = torch.where(yprev > 0, torch.tensor(2.0), torch.where(yprev < 0, torch.tensor(0.2), torch.tensor(0.0)))
        wpd = ops.pack_conv_weight(ops.CONV_K3S1, wt.cuda(), n_is_d3=False, flip=True, tf32=True)
        dx = ops.conv_igemm(ops.CONV_K3S1, dy.cuda(), wpd, yprev=yprev.cuda(), dact=ops.ACT_LEAKY, dact_alpha=0.1,
                            dact_scale=2.0, dact_zero=0.0, tf32=True)
        torch.cuda.synchronize()
        assert _rel_l2(dx.cpu(), xt.grad * slope) < TOL


@pytest.mark.parametrize("n,h,w,cin,cout", CASES)
def test_tf32_k4s2_fprop_and_convt_dgrad(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _rand((n, 2 * h, 2 * w, cin), 7)
    wt = _rand((4, 4, cin, cout), 8, 0.05)
    b = _rand((cout,), 9, 0.1)
    wp = ops.pack_conv_weight(ops.CONV_K4S2, wt.cuda(), n_is_d3=True, tf32=True)
    seed = 0x5EED_0F_7F32
    y = ops.conv_igemm(ops.CONV_K4S2, x.cuda(), wp, bias=b.cuda(), act=ops.ACT_LEAKY, alpha=0.1, dropout_thresh8=128,
                       dropout_seed=seed, tf32=True)
    torch.cuda.synchronize()
    pre = K.leaky_relu(K.conv2d_same(x, wt, b, 2))
    mask = torch.from_numpy(keep_mask(pre.numel(), seed, 128).astype(np.float32)).reshape(pre.shape)
    assert _rel_l2(y.cpu(), K.dropout(pre, mask, 0.5)) < TOL
    # the input gradient of the stride-2 conv is the convT role
    if cin % 64:
        return
    dy = _rand((n, h, w, cout), 10)
    xt = torch.zeros((n, 2 * h, 2 * w, cin), requires_grad=True)
    K.conv2d_same(xt, wt, None, 2).backward(dy)
    wpd = ops.pack_conv_weight(ops.CONVT_K4S2, wt.cuda(), n_is_d3=False, tf32=True)
    dx = ops.conv_igemm(ops.CONVT_K4S2, dy.cuda(), wpd, tf32=True)
    torch.cuda.synchronize()
    assert _rel_l2(dx.cpu(), xt.grad) < TOL


@pytest.mark.parametrize("n,h,w,cin,cout", CASES[:4])
def test_tf32_convt_fprop(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _rand((n, h, w, cin), 11)
    wt = _rand((4, 4, cout, cin), 12, 0.05)
    beta = _rand((cout,), 13, 0.1)
    wp = ops.pack_conv_weight(ops.CONVT_K4S2, wt.cuda(), n_is_d3=False, tf32=True)
    y = ops.conv_igemm(ops.CONVT_K4S2, x.cuda(), wp, bias=beta.cuda(), act=ops.ACT_RELU, tf32=True)
    torch.cuda.synchronize()
    assert _rel_l2(y.cpu(), torch.relu(K.conv2d_transpose_same(x, wt, 2) + beta)) < TOL


WGRAD_CASES = [
    # (N, H(grid), W(grid), Cin, Cout): 64x64 (zero-filled M half), transposed store (Cin=64), N tiles of 64 and 128
    (2, 8, 16, 64, 64),
    (2, 8, 16, 64, 128),
    (3, 18, 32, 128, 128),
    (2, 36, 64, 128, 256),
    (1, 72, 128, 128, 64),
    (2, 18, 32, 256, 512),
]


@pytest.mark.parametrize("n,h,w,cin,cout", WGRAD_CASES)
def test_tf32_wgrad_k3s1(n, h, w, cin, cout):
    """Weight gradient of a stride-1 3x3 conv from UNROUNDED fp32 operands (kind::tf32, MN-major) at 1e-3."""
    from imagegeneration_b200 import ops
    x = _rand((n, h, w, cin), 21)
    dy = _rand((n, h, w, cout), 22)
    wt = torch.zeros((3, 3, cin, cout), requires_grad=True)
    K.conv2d_same(x, wt, None, 1).backward(dy)
    dw = torch.zeros((3, 3, cin, cout), device="cuda")
    ops.conv_wgrad(ops.CONV_K3S1, x.cuda(), dy.cuda(), dw, tf32=True)
    torch.cuda.synchronize()
    assert _rel_l2(dw.cpu(), wt.grad) < TOL


@pytest.mark.parametrize("n,h,w,cin,cout", WGRAD_CASES)
def test_tf32_wgrad_k4s2_conv_and_convt(n, h, w, cin, cout):
    from imagegeneration_b200 import ops
    x = _rand((n, 2 * h, 2 * w, cin), 23)
    dy = _rand((n, h, w, cout), 24)
    wt = torch.zeros((4, 4, cin, cout), requires_grad=True)
    K.conv2d_same(x, wt, None, 2).backward(dy)
    dw = torch.zeros((4, 4, cin, cout), device="cuda")
    ops.conv_wgrad(ops.CONV_K4S2, x.cuda(), dy.cuda(), dw, tf32=True)
    torch.cuda.synchronize()
    assert _rel_l2(dw.cpu(), wt.grad) < TOL
    # Conv2DTranspose kernel [4,4,Cout,Cin]: shifted operand = dY, unshifted = the layer input
    xt_in = _rand((n, h, w, cin), 25)
    dyt = _rand((n, 2 * h, 2 * w, cout), 26)
    wtt = torch.zeros((4, 4, cout, cin), requires_grad=True)
    K.conv2d_transpose_same(xt_in, wtt, 2).backward(dyt)
    dwt = torch.zeros((4, 4, cout, cin), device="cuda")
    ops.conv_wgrad(ops.CONV_K4S2, dyt.cuda(), xt_in.cuda(), dwt, tf32=True)
    torch.cuda.synchronize()
    assert _rel_l2(dwt.cpu(), wtt.grad) < TOL


def test_tf32_wgrad_rejects_bf16_operands():
    from imagegeneration_b200 import ops
    from imagegeneration_b200.lib import B200GanError
    x = torch.zeros((1, 8, 16, 64), device="cuda", dtype=torch.bfloat16)
    dw = torch.zeros((3, 3, 64, 64), device="cuda")
    with pytest.raises(B200GanError):
        ops.conv_wgrad(ops.CONV_K3S1, x, x, dw, tf32=True)
