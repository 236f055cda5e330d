"""PyTorch custom ops (`torch.library`, namespace `b200gan`) with autograd formulas over the C ABI.

north_star: "Python host code drives PyTorch custom ops over a thin C-ABI layer".  The training step itself
(`SNDCGANEngine.train_step`) schedules its launches by hand — backward order, fused derivative epilogues, deferred Adam —
because that is what makes it fast.  This module is the COMPOSABLE form of the same kernels: every layer family of
sndcgan/SNDCGAN.py:25-128 is one differentiable op, so `disc_model(x)` / `gen_model(z)` build an ordinary autograd graph
and a maintainer can put any loss on top and call `.backward()` (see Generator.__call__ / Discriminator.__call__ and
`parameters()` in sndcgan.py).  The forward launches are identical to the engine's; a backward formula differs from the
engine's fused schedule in one respect only: autograd hands a layer the gradient of its OUTPUT, so the activation /
dropout derivative runs as a stand-alone pass (b2g_dact) instead of inside the upstream input-gradient epilogue.

    b200gan::conv2d                 Conv2D k3s1 / k4s2 SAME + bias + LeakyReLU + Dropout        SNDCGAN.py:73-120
    b200gan::conv2d_transpose_bn    Conv2DTranspose k4s2 SAME + inference BatchNorm + ReLU      SNDCGAN.py:36-55
    b200gan::conv2d_transpose_tanh  Conv2DTranspose k3s1 SAME 64 -> 3 + tanh                    SNDCGAN.py:57-61
    b200gan::dense_bn_relu          Dense(128 -> F, no bias) + inference BatchNorm + ReLU       SNDCGAN.py:31-34
    b200gan::dense_head             Flatten + Dense(1)                                          SNDCGAN.py:122-123
    b200gan::bce_with_logits        BinaryCrossentropy(from_logits=True)                        SNDCGAN.py:170

All tensors are CUDA tensors; there is no CPU kernel registered (a CPU call fails: no fallback).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import torch
from torch import Tensor

from . import lib as _lib
from . import ops
from .lib import (ACT_LEAKY, ACT_NONE, ACT_RELU, ACT_TANH, CONV_K1S1, CONV_K3S1, CONV_K3S1_C3, CONV_K4S2, CONVT_K4S2,
                  B200GanError, check)

BN_EPS = 1e-3
INV_STD = 1.0 / math.sqrt(1.0 + BN_EPS)
_U64 = 0xFFFFFFFFFFFFFFFF


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _to_i64(seed: int) -> int:
    seed &= _U64
    return seed - (1 << 64) if seed >= (1 << 63) else seed


def _dact(g: Tensor, y: Tensor, kind: int, alpha: float = 0.0, scale: float = 1.0, zero: float = 0.0) -> Tensor:
    """g * act'(y) from the saved output (b2g_dact)."""
    g = g.contiguous()
    if g.dtype != y.dtype or g.shape != y.shape:
        raise B200GanError("dact: gradient and saved output must match")
    out = torch.empty_like(g)
    check(_lib.load().b2g_dact(_p(g), _p(y), _p(out), g.numel(), int(g.dtype == torch.float32), kind, alpha, scale, zero,
                               _stream()), "b2g_dact")
    return out


def _colsum(x: Tensor, cout: int) -> Tensor:
    out = torch.zeros(cout, dtype=torch.float32, device=x.device)
    check(_lib.load().b2g_colsum_bf16(_p(x), _p(out), x.numel() // cout, cout, _stream()), "b2g_colsum_bf16")
    return out


# ----------------------------------------------------------------------------------------------------------
# Conv2D + bias + LeakyReLU + Dropout
# ----------------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200gan::conv2d", mutates_args=())
def conv2d(x: Tensor, w: Tensor, bias: Optional[Tensor], stride: int, alpha: float, drop_thresh8: int,
           seed: int) -> Tensor:
    """x: bf16 [N,H,W,Cin] (fp32 when Cin == 3), w: fp32 [k,k,Cin,Cout] (k3 s1 or k4 s2, SAME), bias fp32 [Cout].
    LeakyReLU(alpha) then dropout with rate drop_thresh8/256 (0: none) keyed by `seed`.  Returns bf16 NHWC."""
    k, cin = w.shape[0], w.shape[2]
    seed &= _U64
    if cin == 3:
        wp = ops.pack_conv_weight(CONV_K1S1, w.contiguous().view(1, 1, 27, w.shape[3]), n_is_d3=True, ck_pad=64)
        return ops.conv_igemm(CONV_K3S1_C3, x.contiguous(), wp, bias=bias, act=ACT_LEAKY, alpha=alpha,
                              dropout_thresh8=drop_thresh8, dropout_seed=seed)
    kind = CONV_K3S1 if (k, stride) == (3, 1) else CONV_K4S2
    if (k, stride) not in ((3, 1), (4, 2)):
        raise B200GanError("b200gan::conv2d supports k3 s1 and k4 s2")
    wp = ops.pack_conv_weight(kind, w.contiguous(), n_is_d3=True)
    return ops.conv_igemm(kind, x.contiguous(), wp, bias=bias, act=ACT_LEAKY, alpha=alpha,
                          dropout_thresh8=drop_thresh8, dropout_seed=seed)


@conv2d.register_fake
def _(x, w, bias, stride, alpha, drop_thresh8, seed):
    n, h, wd, _ = x.shape
    return x.new_empty((n, h // stride, wd // stride, w.shape[3]), dtype=torch.bfloat16)


def _conv2d_setup(ctx, inputs, output):
    x, w, bias, stride, alpha, drop_thresh8, _seed = inputs
    ctx.save_for_backward(x, w, output)
    ctx.stride, ctx.alpha, ctx.t8, ctx.has_bias = stride, alpha, drop_thresh8, bias is not None


def _conv2d_backward(ctx, gy):
    x, w, y = ctx.saved_tensors
    k, cin, cout = w.shape[0], w.shape[2], w.shape[3]
    scale = 256.0 / (256.0 - ctx.t8) if ctx.t8 else 1.0
    zero = 0.0 if ctx.t8 else ctx.alpha        # y == 0: dropped (training) / LeakyReLU'(0) = alpha (Keras: x > 0 ? 1 : alpha)
    dpre = _dact(gy, y, ACT_LEAKY, ctx.alpha, scale, zero)
    dw = torch.zeros_like(w)
    dx = db = None
    if cin == 3:
        db = torch.zeros(cout, dtype=torch.float32, device=w.device)
        ops.conv_wgrad(CONV_K3S1_C3, x, dpre, dw, dbias=db)
        if ctx.needs_input_grad[0]:
            wpd = ops.pack_conv_weight(CONV_K3S1, w.contiguous(), n_is_d3=False, flip=True)
            dx = ops.conv_igemm(CONV_K3S1, dpre, wpd, out_f32=True)
    else:
        kind = CONV_K3S1 if ctx.stride == 1 else CONV_K4S2
        ops.conv_wgrad(kind, x, dpre, dw)
        db = _colsum(dpre, cout)
        if ctx.needs_input_grad[0]:
            if ctx.stride == 1:
                wpd = ops.pack_conv_weight(CONV_K3S1, w.contiguous(), n_is_d3=False, flip=True)
                dx = ops.conv_igemm(CONV_K3S1, dpre, wpd)
            else:
                wpd = ops.pack_conv_weight(CONVT_K4S2, w.contiguous(), n_is_d3=False)
                dx = ops.conv_igemm(CONVT_K4S2, dpre, wpd)
    return dx, dw, (db if ctx.has_bias else None), None, None, None, None


conv2d.register_autograd(_conv2d_backward, setup_context=_conv2d_setup)


# ----------------------------------------------------------------------------------------------------------
# Conv2DTranspose k4 s2 (no bias) + BatchNorm (inference statistics 0 / 1) + ReLU
# ----------------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200gan::conv2d_transpose_bn", mutates_args=())
def conv2d_transpose_bn(x: Tensor, w: Tensor, gamma: Tensor, beta: Tensor) -> Tensor:
    """x bf16 [N,H,W,Cin], w fp32 [4,4,Cout,Cin] (Keras Conv2DTranspose layout), gamma/beta fp32 [Cout]:
    relu(gamma * convT(x, w) / sqrt(1 + eps) + beta), bf16 [N,2H,2W,Cout].  gamma/sqrt(1+eps) is folded into the packed
    weight exactly like the engine does."""
    inv = torch.full((1,), INV_STD, dtype=torch.float32, device=x.device)
    wp = ops.pack_conv_weight(CONVT_K4S2, w.contiguous(), n_is_d3=False, scale_scalar=inv, scale_d2=gamma)
    return ops.conv_igemm(CONVT_K4S2, x.contiguous(), wp, bias=beta, act=ACT_RELU)


@conv2d_transpose_bn.register_fake
def _(x, w, gamma, beta):
    n, h, wd, _ = x.shape
    return x.new_empty((n, 2 * h, 2 * wd, w.shape[2]), dtype=torch.bfloat16)


def _convt_setup(ctx, inputs, output):
    x, w, gamma, _beta = inputs
    ctx.save_for_backward(x, w, gamma, output)


def _convt_backward(ctx, gy):
    x, w, gamma, y = ctx.saved_tensors
    cout, cin = w.shape[2], w.shape[3]
    lib = _lib.load()
    dpre = _dact(gy, y, ACT_RELU)
    dbeta = _colsum(dpre, cout)
    dw = torch.zeros_like(w)
    ops.conv_wgrad(CONV_K4S2, dpre, x, dw)                    # raw gradient w.r.t. the folded weight
    dgamma = torch.empty(cout, dtype=torch.float32, device=w.device)
    check(lib.b2g_bn_fold_finalize(_p(w), _p(dw), _p(gamma), INV_STD, _p(dgamma), 16, cout, cin, _stream()),
          "b2g_bn_fold_finalize")
    dx = None
    if ctx.needs_input_grad[0]:
        inv = torch.full((1,), INV_STD, dtype=torch.float32, device=x.device)
        wpd = ops.pack_conv_weight(CONV_K4S2, w.contiguous(), n_is_d3=True, scale_scalar=inv, scale_d2=gamma)
        dx = ops.conv_igemm(CONV_K4S2, dpre, wpd)
    return dx, dw, dgamma, dbeta


conv2d_transpose_bn.register_autograd(_convt_backward, setup_context=_convt_setup)


# ----------------------------------------------------------------------------------------------------------
# Conv2DTranspose k3 s1 64 -> 3 + tanh (fp32 image)
# ----------------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200gan::conv2d_transpose_tanh", mutates_args=())
def conv2d_transpose_tanh(x: Tensor, w: Tensor) -> Tensor:
    """x bf16 [N,H,W,64], w fp32 [3,3,3,64] -> tanh(convT(x, w)) fp32 [N,H,W,3]."""
    wp = ops.pack_conv_weight(CONV_K3S1, w.contiguous(), n_is_d3=False, flip=True)
    return ops.conv_igemm(CONV_K3S1, x.contiguous(), wp, act=ACT_TANH, out_f32=True)


@conv2d_transpose_tanh.register_fake
def _(x, w):
    n, h, wd, _ = x.shape
    return x.new_empty((n, h, wd, 3), dtype=torch.float32)


def _convt_tanh_setup(ctx, inputs, output):
    x, w = inputs
    ctx.save_for_backward(x, w, output)


def _convt_tanh_backward(ctx, gy):
    x, w, y = ctx.saved_tensors
    dpre = _dact(gy.to(torch.float32), y, ACT_TANH)
    dw = torch.zeros_like(w)
    ops.conv_wgrad(CONV_K3S1_C3, dpre, x, dw)
    dx = None
    if ctx.needs_input_grad[0]:
        wpd = ops.pack_conv_weight(CONV_K1S1, w.contiguous().view(1, 1, 27, 64), n_is_d3=True, ck_pad=64)
        dx = ops.conv_igemm(CONV_K3S1_C3, dpre, wpd)
    return dx, dw


conv2d_transpose_tanh.register_autograd(_convt_tanh_backward, setup_context=_convt_tanh_setup)


# ----------------------------------------------------------------------------------------------------------
# Dense(128 -> F) + BatchNorm(inference) + ReLU   (generator stem)
# ----------------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200gan::dense_bn_relu", mutates_args=())
def dense_bn_relu(z: Tensor, w: Tensor, gamma: Tensor, beta: Tensor) -> Tensor:
    """z fp32 [B,128], w fp32 [128,F], gamma/beta fp32 [F] -> relu(gamma * (z w) / sqrt(1 + eps) + beta), bf16 [B,F]."""
    B, F = z.shape[0], w.shape[1]
    out = torch.empty((B, F), dtype=torch.bfloat16, device=z.device)
    check(_lib.load().b2g_dense_g_fwd(_p(z.contiguous()), _p(w), _p(gamma), _p(beta), INV_STD, _p(out), B, F, _stream()),
          "b2g_dense_g_fwd")
    return out


@dense_bn_relu.register_fake
def _(z, w, gamma, beta):
    return z.new_empty((z.shape[0], w.shape[1]), dtype=torch.bfloat16)


def _stem_setup(ctx, inputs, output):
    z, w, gamma, _beta = inputs
    ctx.save_for_backward(z, w, gamma, output)


def _stem_backward(ctx, gy):
    z, w, gamma, y = ctx.saved_tensors
    B, F = z.shape[0], w.shape[1]
    dpre = _dact(gy, y, ACT_RELU)
    dw, dgamma, dbeta = torch.empty_like(w), torch.empty_like(gamma), torch.empty_like(gamma)
    check(_lib.load().b2g_dense_g_bwd(_p(z.contiguous()), _p(w), _p(gamma), INV_STD, _p(dpre), _p(dw), _p(dgamma),
                                      _p(dbeta), B, F, _stream()), "b2g_dense_g_bwd", (B + 31) // 32)
    return None, dw, dgamma, dbeta            # the latent vector is an input, not a variable (SNDCGAN.py:248)


dense_bn_relu.register_autograd(_stem_backward, setup_context=_stem_setup)


# ----------------------------------------------------------------------------------------------------------
# Flatten + Dense(1)   (discriminator head)
# ----------------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200gan::dense_head", mutates_args=())
def dense_head(x: Tensor, w: Tensor, bias: Tensor) -> Tensor:
    """x bf16 [B, ...] (NHWC, flattened row-major like Keras Flatten), w fp32 [F,1], bias fp32 [1] -> logits fp32 [B,1]."""
    B = x.shape[0]
    F = x.numel() // B
    dev = x.device
    ws = torch.empty(64 * B, dtype=torch.float32, device=dev)
    logits = torch.empty((B, 1), dtype=torch.float32, device=dev)
    scratch = torch.empty(B + 2, dtype=torch.float32, device=dev)
    check(_lib.load().b2g_dhead_fwd_loss(_p(x.contiguous()), _p(w), None, _p(bias), B, F, 0, 1.0, 1.0 / B, _p(ws),
                                         ws.numel(), _p(logits), _p(scratch), _p(scratch[B:]), _stream()),
          "b2g_dhead_fwd_loss", 2)
    return logits


@dense_head.register_fake
def _(x, w, bias):
    return x.new_empty((x.shape[0], 1), dtype=torch.float32)


def _head_setup(ctx, inputs, output):
    x, w, _bias = inputs
    ctx.save_for_backward(x, w)


def _head_backward(ctx, gy):
    x, w = ctx.saved_tensors
    B = x.shape[0]
    F = x.numel() // B
    dl = gy.to(torch.float32).contiguous().view(B)
    dx = torch.empty_like(x)
    dw, db = torch.zeros_like(w), torch.zeros(1, dtype=torch.float32, device=w.device)
    # alpha = dscale = dzero = 1: no activation derivative here (the conv op in front applies its own)
    check(_lib.load().b2g_dhead_bwd(_p(x), _p(w), None, _p(dl), B, F, 1.0, 1.0, 1.0, _p(dx), _p(dw), _p(db), _stream()),
          "b2g_dhead_bwd")
    return dx, dw, db


dense_head.register_autograd(_head_backward, setup_context=_head_setup)


# ----------------------------------------------------------------------------------------------------------
# BinaryCrossentropy(from_logits=True)
# ----------------------------------------------------------------------------------------------------------
@torch.library.custom_op("b200gan::bce_with_logits", mutates_args=())
def bce_with_logits(y_true: Tensor, y_pred: Tensor) -> Tensor:
    """mean_b[max(x,0) - x*y + log1p(exp(-|x|))] over the batch (Keras reduction SUM_OVER_BATCH_SIZE); fp32 scalar."""
    B = y_pred.numel()
    out = torch.empty((), dtype=torch.float32, device=y_pred.device)
    check(_lib.load().b2g_bce_logits(_p(y_pred.to(torch.float32).contiguous()), _p(y_true.to(torch.float32).contiguous()),
                                     B, _p(out), None, _stream()), "b2g_bce_logits")
    return out


@bce_with_logits.register_fake
def _(y_true, y_pred):
    return y_pred.new_empty((), dtype=torch.float32)


def _bce_setup(ctx, inputs, output):
    y_true, y_pred = inputs
    ctx.save_for_backward(y_true, y_pred)


def _bce_backward(ctx, g):
    y_true, y_pred = ctx.saved_tensors
    B = y_pred.numel()
    out = torch.empty((), dtype=torch.float32, device=y_pred.device)
    dl = torch.empty(B, dtype=torch.float32, device=y_pred.device)
    check(_lib.load().b2g_bce_logits(_p(y_pred.to(torch.float32).contiguous()), _p(y_true.to(torch.float32).contiguous()),
                                     B, _p(out), _p(dl), _stream()), "b2g_bce_logits")
    return None, (dl.view_as(y_pred) * g).to(y_pred.dtype)


bce_with_logits.register_autograd(_bce_backward, setup_context=_bce_setup)


class BinaryCrossentropy:
    """tf.keras.losses.BinaryCrossentropy(from_logits=True) — the reference's `loss_fn` attribute (SNDCGAN.py:170):
    `loss_fn(y_true, y_pred)` on CUDA tensors, differentiable w.r.t. y_pred."""

    def __init__(self, from_logits: bool = True):
        if not from_logits:
            raise B200GanError("only from_logits=True (the reference's setting) is implemented")
        self.from_logits = True

    def __call__(self, y_true: Tensor, y_pred: Tensor) -> Tensor:
        if not y_pred.is_cuda:
            raise B200GanError("b200gan ops need CUDA tensors: there is no CPU fallback")
        return bce_with_logits(y_true, y_pred)


# ----------------------------------------------------------------------------------------------------------
# model-level graphs (used by Generator.__call__ / Discriminator.__call__ when autograd is recording)
# ----------------------------------------------------------------------------------------------------------
def generator_graph(gen, z: Tensor) -> Tensor:
    """sndcgan/SNDCGAN.py:28-61 as differentiable ops; `gen.parameters()` are the leaves."""
    p = gen.parameters()
    h0, w0 = gen.h0, gen.w0
    x = dense_bn_relu(z, p[0], p[1], p[2]).view(z.shape[0], h0, w0, 512)
    for k in (3, 6, 9):
        x = conv2d_transpose_bn(x, p[k], p[k + 1], p[k + 2])
    return conv2d_transpose_tanh(x, p[12])


def discriminator_graph(disc, x: Tensor, training: bool, dropout_seeds) -> Tensor:
    """sndcgan/SNDCGAN.py:70-124 as differentiable ops; `disc.parameters()` are the leaves."""
    if disc.spectral_norm:
        raise B200GanError("the autograd composition does not cover spectral_norm=True; use SNDCGANEngine.train_step")
    from .sndcgan import D_CONVS, LEAKY_ALPHA
    p = disc.parameters()
    t8 = disc.thresh8 if training else 0
    seeds = dropout_seeds if t8 else [0] * 7
    if t8 and dropout_seeds is None:
        raise B200GanError("training-mode forward needs dropout_seeds")
    for i, (_name, _k, s, _ci, _co) in enumerate(D_CONVS):
        x = conv2d(x, p[2 * i], p[2 * i + 1], s, LEAKY_ALPHA, t8, _to_i64(seeds[i]))
    return dense_head(x, p[14], p[15])
