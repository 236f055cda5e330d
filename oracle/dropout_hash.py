"""ORACLE — TEST INFRASTRUCTURE ONLY.

Bit-exact numpy restatement of the counter-based dropout RNG in imagegeneration_b200/csrc/b2g_ptx.cuh
(b2g_mix32 / b2g_dropout_bits).  TensorFlow's stateful RNG stream (tf.nn.dropout behind
tf.keras.layers.Dropout, sndcgan/SNDCGAN.py:78) cannot be reproduced outside TF, so parity tests feed the
oracle the very keep-masks the CUDA kernels apply: element i of a flat NHWC activation is DROPPED when
byte (i & 3) of hash(i >> 2, seed) is < thresh8, i.e. with probability thresh8 / 256; rate 0.5 (thresh8 = 128) spends
one bit per element instead (hash(i >> 5, seed), see keep_mask).
"""
from __future__ import annotations

import numpy as np

_M1 = np.uint32(0x7FEB352D)
_M2 = np.uint32(0x846CA68B)


def mix32(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint32, copy=True)
    x ^= x >> np.uint32(16)
    x *= _M1
    x ^= x >> np.uint32(15)
    x *= _M2
    x ^= x >> np.uint32(16)
    return x


def dropout_bits(group: np.ndarray, seed: int) -> np.ndarray:
    """b2g_dropout_bits: lowbias32 of (group ^ seed_lo) with seed_hi added between its two multiply rounds."""
    seed_lo = np.uint32(seed & 0xFFFFFFFF)
    seed_hi = np.uint32((seed >> 32) & 0xFFFFFFFF)
    with np.errstate(over="ignore"):
        x = group.astype(np.uint32) ^ seed_lo
        x ^= x >> np.uint32(16)
        x *= _M1
        x ^= x >> np.uint32(15)
        x += seed_hi
        x *= _M2
        x ^= x >> np.uint32(16)
        return x


def keep_mask(numel: int, seed: int, thresh8: int) -> np.ndarray:
    """uint8 keep mask (1 = kept) for a flat activation of `numel` elements (numel % 4 == 0).

    thresh8 == 128 (rate 0.5, the reference's default): one BIT per element — element e = i & 31 of the hash word of
    group i >> 5 is kept iff bit 8 * (e & 3) + 7 - (e >> 2) is set (b2g_keep128 in b2g_ptx.cuh).
    Other rates: one BYTE per element — byte i & 3 of the hash word of group i >> 2 is >= thresh8."""
    assert numel % 4 == 0
    if thresh8 <= 0:
        return np.ones(numel, dtype=np.uint8)
    if thresh8 == 128:
        words = (numel + 31) // 32
        with np.errstate(over="ignore"):
            bits = dropout_bits(np.arange(words, dtype=np.uint32), seed)
        e = np.arange(32, dtype=np.uint32)
        pos = np.uint32(8) * (e & np.uint32(3)) + np.uint32(7) - (e >> np.uint32(2))
        keep = (bits[:, None] >> pos[None, :]) & np.uint32(1)
        return keep.reshape(-1)[:numel].astype(np.uint8)
    groups = np.arange(numel // 4, dtype=np.uint32)
    with np.errstate(over="ignore"):
        bits = dropout_bits(groups, seed)
    bytes_ = np.stack([(bits >> np.uint32(8 * b)) & np.uint32(0xFF) for b in range(4)], axis=1).reshape(-1)
    return (bytes_ >= np.uint32(thresh8)).astype(np.uint8)


def rate_to_thresh8(rate: float) -> int:
    """Dropout rate -> 8-bit threshold (rate quantised to multiples of 1/256; 0.5 is exact)."""
    t = int(round(rate * 256.0))
    return max(0, min(255, t))
