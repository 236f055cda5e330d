"""Per-launch CUDA-event timing of the tensor-core kernels inside one train_step (run on the GPU box)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200 import lib
from imagegeneration_b200.sndcgan import SNDCGANEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
eng = SNDCGANEngine(B, dropout=0.5)
real = torch.rand((B, 144, 256, 3), device="cuda") * 2 - 1
z = torch.rand((B, 128), device="cuda") * 2 - 1
for _ in range(3):
    eng.train_step(real, z)
torch.cuda.synchronize()
lib.profile_sink = []
eng.train_step(real, z)
torch.cuda.synchronize()
tot = 0.0
for name, flops, a, b, label in lib.profile_sink:
    ms = a.elapsed_time(b)
    tot += ms
    print(f"{name:11s} {label:42s} {ms*1e3:8.1f} us  {flops/ms/1e9:7.1f} TF/s")
print("total tensor-core ms", tot)
