"""ORACLE — TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's FID-like score (sndcgan/generator_evaluation.py), SURVEY.md section 8(f) row N3:
  calculate_fid          generator_evaluation.py:36-55   (numpy mean / cov(rowvar=False) / scipy.linalg.sqrtm)
  feature extractor      generator_evaluation.py:134-138 (discriminator minus Flatten and Dense — its last layer is then
                         conv7's Dropout, the identity under predict — plus AveragePooling2D(8, 8) and Flatten)
numpy and scipy are the reference's own dependencies for this function and are installed here, so calculate_fid below
is the reference's arithmetic call for call; only the Keras feature extractor is restated (PARITY UNPINNED there).
"""
from __future__ import annotations

import numpy as np
import torch
from scipy.linalg import sqrtm

from . import sndcgan_oracle as O


def avg_pool_8x8(x: torch.Tensor) -> torch.Tensor:
    """tf.keras.layers.AveragePooling2D(pool_size=(8, 8)): stride 8, padding 'valid' (trailing rows/columns that do not
    fill a window are dropped: 18 x 32 -> 2 x 4).  x NHWC."""
    n, h, w, c = x.shape
    ph, pw = h // 8, w // 8
    return x[:, :ph * 8, :pw * 8, :].reshape(n, ph, 8, pw, 8, c).mean(dim=(2, 4))


def features(dv, images: torch.Tensor) -> np.ndarray:
    """disc_model.predict(images) of the popped model (generator_evaluation.py:38-39,134-138): [B, 4096] at 144x256."""
    col = {}
    with torch.no_grad():
        O.discriminator_forward(dv, images, 0.0, None, col)
        f = avg_pool_8x8(col["d_c7"])
    return f.reshape(f.shape[0], -1).numpy()


def calculate_fid_from_features(predict_fake: np.ndarray, predict_real: np.ndarray) -> float:
    """generator_evaluation.py:41-55, verbatim arithmetic."""
    mu_fake = np.mean(predict_fake, axis=0)
    mu_real = np.mean(predict_real, axis=0)
    cov_fake = np.cov(predict_fake, rowvar=False)
    cov_real = np.cov(predict_real, rowvar=False)
    ssdiff = np.sum((mu_fake - mu_real) ** 2.0)
    covmean = sqrtm(np.dot(cov_fake, cov_real))
    if np.iscomplexobj(covmean):
        covmean = covmean.real
    return float(ssdiff + np.trace(cov_fake + cov_real - 2.0 * covmean))


def calculate_fid(dv, images_fake: torch.Tensor, images_real: torch.Tensor) -> float:
    return calculate_fid_from_features(features(dv, images_fake), features(dv, images_real))


def trace_sqrt_product_lowrank(predict_fake: np.ndarray, predict_real: np.ndarray) -> float:
    """tr sqrtm(cov_fake cov_real) without forming the d x d matrices: with centred X (n x d), Y (m x d) the non-zero
    eigenvalues of cov_fake cov_real are the squared singular values of X Y^T / sqrt((n-1)(m-1)), so the trace of the
    principal square root is the nuclear norm of that n x m matrix.  (What the CUDA path evaluates; exact in exact
    arithmetic, and better conditioned than sqrtm of a singular 4096 x 4096 product.)"""
    x = predict_fake - predict_fake.mean(axis=0)
    y = predict_real - predict_real.mean(axis=0)
    m = x.astype(np.float64) @ y.astype(np.float64).T / np.sqrt((x.shape[0] - 1) * (y.shape[0] - 1))
    return float(np.linalg.svd(m, compute_uv=False).sum())
