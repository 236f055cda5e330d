"""Regenerates tests/golden/oracle_tiny.npz from the CPU oracle (oracle/sndcgan_oracle.py).

The reference ships no golden vectors and TensorFlow cannot run here, so these fixtures do NOT pin
the oracle to TensorFlow ("parity unpinned"); they freeze the oracle's own outputs so that any later change to the
restatement is caught, and they carry the known-answer vector of the dropout hash shared with the CUDA kernels.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import sndcgan_oracle as O  # noqa: E402
from oracle.dropout_hash import dropout_bits, keep_mask  # noqa: E402

IMAGE = (16, 32, 3)
B = 2


def inputs():
    g = torch.Generator().manual_seed(2024)
    real = torch.rand((B, *IMAGE), generator=g, dtype=torch.float64) * 2 - 1
    z = torch.rand((B, 128), generator=g, dtype=torch.float64) * 2 - 1
    return real, z


def masks(pass_idx):
    h, w, _ = IMAGE
    shapes = [(h, w, 64), (h // 2, w // 2, 128), (h // 2, w // 2, 128), (h // 4, w // 4, 256), (h // 4, w // 4, 256),
              (h // 8, w // 8, 512), (h // 8, w // 8, 512)]
    return [torch.from_numpy(keep_mask(B * a * b * c, 1000 * (pass_idx + 1) + i, 128).astype(np.float64)).reshape(B, a, b, c)
            for i, (a, b, c) in enumerate(shapes)]


def run():
    orc = O.OracleSNDCGAN(IMAGE, dropout=0.5, seed=62, dtype=torch.float64)
    real, z = inputs()
    out = {}
    for step in range(2):
        col = {}
        r = orc.train_step(real, z, masks(0), masks(1), masks(2), col)
        out[f"losses{step}"] = np.array([float(x) for x in r[:4]])
        out[f"logits_real{step}"] = r[4].numpy()
        out[f"logits_fake{step}"] = r[5].numpy()
        out[f"g_grad_norms{step}"] = np.array([float(g.norm()) for g in col["g_grads"]])
        out[f"d_grad_norms{step}"] = np.array([float(g.norm()) for g in col["d_grads2"]])
    out["g_var_sums"] = np.array([float(v.sum()) for v in orc.gv])
    out["d_var_sums"] = np.array([float(v.sum()) for v in orc.dv])
    out["samples_mean"] = np.array([float(O.create_samples(orc.gv, z, B, IMAGE).mean())])
    out["hash_kat_bits"] = dropout_bits(np.arange(8, dtype=np.uint32), 0x123456789ABCDEF1)
    out["hash_kat_mask"] = keep_mask(64, 0x123456789ABCDEF1, 128)
    return out


if __name__ == "__main__":
    np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle_tiny.npz"), **run())
    print("written")
