"""One eager training step of every engine flavour at small batch, for compute-sanitizer (memcheck / racecheck /
synccheck): full-resolution SNDCGAN step at batch 2 (row-slab, CTA-pair, per-tap, 3-channel edge, weight-gradient, stem,
head, Adam and packing kernels), a spectral-norm step (b2g_adam_sn, b2g_sn_dot_multi, power iteration) and a tf32 step.
    compute-sanitizer --tool racecheck python tools/sanitizer_probe.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200.sndcgan import SNDCGANEngine

which = sys.argv[1] if len(sys.argv) > 1 else "all"
g = torch.Generator().manual_seed(1)


def step(eng, size, batch):
    real = (torch.rand((batch, *size), generator=g) * 2 - 1).cuda()
    z = (torch.rand((batch, 128), generator=g) * 2 - 1).cuda()
    out = eng.train_step(real, z)
    torch.cuda.synchronize()
    print([round(float(x), 5) for x in out[:4]], flush=True)


if which in ("all", "bf16"):
    step(SNDCGANEngine(2, dropout=0.5, image_size=(144, 256, 3), cuda_graph=False), (144, 256, 3), 2)
if which in ("all", "sn"):
    step(SNDCGANEngine(4, dropout=0.5, image_size=(48, 64, 3), cuda_graph=False, spectral_norm=True, loss="hinge"), (48, 64, 3), 4)
if which in ("all", "tf32"):
    step(SNDCGANEngine(2, dropout=0.5, image_size=(48, 64, 3), dtype="tf32"), (48, 64, 3), 2)
print("done")
