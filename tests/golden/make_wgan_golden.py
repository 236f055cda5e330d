"""Regenerates tests/golden/wgan_oracle_tiny.npz from the CPU oracle (oracle/wgan_oracle.py).

The reference ships no golden vectors for wasserstein_gan/ and TensorFlow cannot run here, so this fixture does NOT
pin the oracle to TensorFlow ("parity unpinned"); it freezes the oracle's own outputs (fp64) so that a later change
to the restatement is caught.

    python tests/golden/make_wgan_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import wgan_oracle as W  # noqa: E402

IMAGE = (16, 32, 3)
B = 4


def inputs():
    g = torch.Generator().manual_seed(2025)
    reals = [torch.rand((B, *IMAGE), generator=g, dtype=torch.float64) * 2 - 1 for _ in range(2)]
    zs = [torch.randn((B, 128), generator=g, dtype=torch.float64) for _ in range(3)]
    return reals, zs


def run():
    orc = W.OracleWGAN(IMAGE, seed=62, dtype=torch.float64)
    reals, zs = inputs()
    out = {}
    for it in range(2):
        c1, c2, gl = orc.train_iteration(reals, zs[:2], zs[2])
        out[f"losses{it}"] = np.array([float(c1), float(c2), float(gl)])
    # conv biases are left out: BatchNorm cancels them, so their gradients (and values) are pure rounding noise
    out["c_var_sums"] = np.array([float(v.sum()) for i, v in enumerate(orc.cv) if i >= 28 or i % 4 != 1])
    out["g_var_sums"] = np.array([float(v.sum()) for v in orc.gv])
    out["c_moving_var_means"] = np.array([float(m[1].mean()) for m in orc.c_moving])
    out["g_moving_mean_sums"] = np.array([float(m[0].sum()) for m in orc.g_moving])
    out["c_kernel_absmax"] = np.array([float(orc.cv[4 * i].abs().max()) for i in range(7)])
    out["sample_mean"] = np.array([float(orc.predict(zs[0]).mean())])
    return out


if __name__ == "__main__":
    np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "wgan_oracle_tiny.npz"), **run())
    print("written")
