"""Regenerates tests/golden/bench_step0.json: the CPU oracle's outputs (plain fp32, oracle/sndcgan_oracle.py) for the
FIRST train_step of exactly what `bench.py` runs — engine seed 62, rank 0, bench.py's synthetic batch (torch CPU
generator seed 1234), dropout 0.5 with the engine's own step-0 dropout seeds (masks regenerated with
oracle/dropout_hash.py) — at batch 16 (BASELINE configs[0]) and batch 64 (configs[1], the benched dispatch).

bench.py compares its first warm-up step with these values in the same process that is timed, and
tests/test_baseline_configs_gpu.py checks the same numbers.  The oracle itself stays "parity unpinned" (no TensorFlow
here, no vectors in the reference): this file pins the CUDA path to the oracle at the BASELINE configs, nothing more.

    python tests/golden/make_bench_golden.py            # ~2 min on 8 cores, ~12 GB of RAM at batch 64
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from imagegeneration_b200.sndcgan import rate_to_thresh8, step_dropout_seeds  # noqa: E402  (pure Python, no CUDA)
from oracle import sndcgan_oracle as O  # noqa: E402
from oracle.dropout_hash import keep_mask  # noqa: E402

IMAGE = (144, 256, 3)
SEED = 62


def bench_inputs(batch, rank=0):
    """The first of bench.py's two synthetic host batches (same generator call order as bench.run_b200)."""
    g = torch.Generator().manual_seed(1234 + rank)
    real = [torch.rand((batch, *IMAGE), generator=g) * 2 - 1 for _ in range(2)]
    z = [torch.rand((batch, 128), generator=g) * 2 - 1 for _ in range(2)]
    return real[0], z[0]


def masks(batch, pass_idx, step=0, rank=0, rate=0.5):
    h, w, _ = IMAGE
    shapes = [(h, w, 64), (h // 2, w // 2, 128), (h // 2, w // 2, 128), (h // 4, w // 4, 256), (h // 4, w // 4, 256),
              (h // 8, w // 8, 512), (h // 8, w // 8, 512)]
    seeds = step_dropout_seeds(SEED, rank, step, pass_idx)
    t8 = rate_to_thresh8(rate)
    return [torch.from_numpy(keep_mask(batch * a * b * c, s, t8).astype(np.float32)).reshape(batch, a, b, c)
            for s, (a, b, c) in zip(seeds, shapes)]


def oracle_step0(batch):
    orc = O.OracleSNDCGAN(IMAGE, dropout=0.5, seed=SEED)      # same initial weights as SNDCGANEngine(seed=62)
    real, z = bench_inputs(batch)
    r = orc.train_step(real, z, masks(batch, 0), masks(batch, 1), masks(batch, 2))
    return {"losses": [float(x) for x in r[:4]],
            "d_logits_real": [float(x) for x in r[4].flatten()],
            "d_logits_fake": [float(x) for x in r[5].flatten()]}


if __name__ == "__main__":
    out = {"about": "oracle (fp32 CPU restatement) outputs of bench.py's first train_step; see make_bench_golden.py",
           "seed": SEED, "image_size": list(IMAGE), "dropout": 0.5, "torch": torch.__version__}
    for b in (16, 64):
        out[f"batch{b}"] = oracle_step0(b)
        print(b, out[f"batch{b}"]["losses"], flush=True)
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "bench_step0.json"), "w") as f:
        json.dump(out, f, indent=0)
    print("written")
