"""ORACLE — TEST INFRASTRUCTURE ONLY (imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline).

CPU restatement, in plain PyTorch (fp32 or fp64), of the TensorFlow/Keras-2 layer arithmetic that the
reference's SNDCGAN hot path executes.  TensorFlow is NOT vendored under /root/reference and cannot be
installed here, so these semantics come from the Keras documentation; the reference ships no tests, golden
vectors or saved weights for this path ==> PARITY UNPINNED at the TensorFlow boundary.

All activations are NHWC; weights use the reference's Keras layouts:
  Conv2D kernel           [kh, kw, Cin, Cout]   (sndcgan/SNDCGAN.py:73-120)
  Conv2DTranspose kernel  [kh, kw, Cout, Cin]   (sndcgan/SNDCGAN.py:36-61)
  Dense kernel            [in, out]             (sndcgan/SNDCGAN.py:31,123)
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

BN_EPS = 1e-3          # tf.keras.layers.BatchNormalization default epsilon (SNDCGAN.py:32,40,47,54)
LEAKY_ALPHA = 0.1      # LeakyReLU(alpha=0.1) (SNDCGAN.py:77)
ADAM_BETA1, ADAM_BETA2, ADAM_EPS = 0.9, 0.999, 1e-7   # tf.keras.optimizers.Adam defaults (SNDCGAN.py:165,168)


def _same_pad(k: int, s: int) -> int:
    # TF 'SAME' for even input sizes: total pad = k - s, split evenly for (k=3,s=1) and (k=4,s=2).
    total = k - s
    assert total % 2 == 0, "asymmetric SAME padding is not needed on the SNDCGAN path"
    return total // 2


def conv2d_same(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor | None, stride: int) -> torch.Tensor:
    """tf.keras.layers.Conv2D(padding='same'): x NHWC, w [kh,kw,Cin,Cout], b [Cout] -> NHWC."""
    kh = w.shape[0]
    y = F.conv2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), b, stride=stride, padding=_same_pad(kh, stride))
    return y.permute(0, 2, 3, 1)


def conv2d_transpose_same(x: torch.Tensor, w: torch.Tensor, stride: int) -> torch.Tensor:
    """tf.keras.layers.Conv2DTranspose(padding='same', use_bias=False): x NHWC, w [kh,kw,Cout,Cin] -> NHWC.

    TF defines it as the gradient of Conv2D w.r.t. its input, no spatial flip; torch's conv_transpose2d has
    the same definition with weight [Cin, Cout, kh, kw] and output size (in-1)*s - 2p + k = in*s."""
    kh = w.shape[0]
    y = F.conv_transpose2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), None, stride=stride,
                           padding=_same_pad(kh, stride))
    return y.permute(0, 2, 3, 1)


def dense(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor | None = None) -> torch.Tensor:
    y = x @ w
    return y if b is None else y + b


def batchnorm_inference(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
                        moving_mean: torch.Tensor | None = None, moving_var: torch.Tensor | None = None,
                        eps: float = BN_EPS) -> torch.Tensor:
    """BatchNormalization(axis=-1) with training=False: y = gamma*(x-mean)/sqrt(var+eps)+beta.

    In the reference's train_step the generator is called without training=True (SNDCGAN.py:243), so the
    moving statistics stay at their initial values mean=0, var=1 forever."""
    mean = torch.zeros_like(gamma) if moving_mean is None else moving_mean
    var = torch.ones_like(gamma) if moving_var is None else moving_var
    return gamma * (x - mean) / torch.sqrt(var + eps) + beta


def leaky_relu(x: torch.Tensor, alpha: float = LEAKY_ALPHA) -> torch.Tensor:
    return torch.where(x > 0, x, alpha * x)


def dropout(x: torch.Tensor, keep_mask: torch.Tensor | None, rate: float) -> torch.Tensor:
    """tf.keras.layers.Dropout(rate) in training mode with an explicit keep mask (1 = kept)."""
    if keep_mask is None or rate == 0.0:
        return x
    return x * keep_mask.to(x.dtype) / (1.0 - rate)


def bce_with_logits(logits: torch.Tensor, label: float) -> torch.Tensor:
    """tf.keras.losses.BinaryCrossentropy(from_logits=True), reduction SUM_OVER_BATCH_SIZE (SNDCGAN.py:170):
    mean over the last axis, then over the batch, of max(x,0) - x*y + log1p(exp(-|x|))."""
    x = logits
    per = torch.clamp(x, min=0) - x * label + torch.log1p(torch.exp(-torch.abs(x)))
    return per.mean(dim=-1).mean()


def hinge_d_loss(logits: torch.Tensor, label: float) -> torch.Tensor:
    """north-star option (not in the reference): D hinge terms mean(relu(1 - x)) for real / mean(relu(1 + x)) for fake."""
    return torch.relu(1.0 - logits).mean() if label == 1.0 else torch.relu(1.0 + logits).mean()


def adam_step(w: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, t: int, lr: float,
              beta1: float = ADAM_BETA1, beta2: float = ADAM_BETA2, eps: float = ADAM_EPS) -> None:
    """Keras-2 Adam._resource_apply_dense (in place).  t is the 1-based step (iterations + 1);
    epsilon sits OUTSIDE the bias-corrected sqrt:  w -= lr_t * m / (sqrt(v) + eps).
    Keras keeps the hyper-parameters in the variable's dtype (fp32) and forms `1 - beta` there
    (Adam._prepare_local: one_minus_beta_2_t = 1 - beta_2_t): 1 - float32(0.999) = 0.00099998713, 1.3e-5 away from
    0.001 — restated here so that v matches an fp32 implementation to rounding error whatever dtype the oracle runs in."""
    import numpy as np
    lr_t = lr * math.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
    c1 = float(np.float32(1.0) - np.float32(beta1))
    c2 = float(np.float32(1.0) - np.float32(beta2))
    m.add_((g - m) * c1)
    v.add_((g * g - v) * c2)
    w.sub_(lr_t * m / (torch.sqrt(v) + eps))


def glorot_uniform(shape: tuple[int, ...], gen: torch.Generator, dtype=torch.float32) -> torch.Tensor:
    """keras glorot_uniform: U(+-sqrt(6/(fan_in+fan_out))); conv fans include the receptive field."""
    if len(shape) == 2:
        fan_in, fan_out = shape
    else:
        rf = 1
        for d in shape[:-2]:
            rf *= d
        fan_in, fan_out = shape[-2] * rf, shape[-1] * rf
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    return ((torch.rand(shape, generator=gen, dtype=torch.float64) * 2 - 1) * limit).to(dtype)


def spectral_norm_power_iteration(w2d: torch.Tensor, u: torch.Tensor, iters: int = 1, eps: float = 1e-12):
    """Miyato et al. (2018) Alg. 1 on W reshaped to [K, Cout] (Keras `reshape(W, (-1, Cout))`); u has length
    Cout.  v <- W u / ||W u||;  u <- W^T v / ||W^T v||;  sigma = v^T W u.  (Not in the reference: SURVEY F1.)"""
    for _ in range(iters):
        v = w2d @ u
        v = v / (v.norm() + eps)
        u = w2d.t() @ v
        u = u / (u.norm() + eps)
    sigma = v @ (w2d @ u)
    return sigma, u, v
