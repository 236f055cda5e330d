"""One launch of a chosen kernel family at the bench shape, for `ncu --set full` (run on the GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imagegeneration_b200 import ops
B = 64
which = sys.argv[1]
if which == "wgrad_c3":
    x = torch.rand((B, 144, 256, 3), device="cuda") * 2 - 1
    dy = torch.rand((B, 144, 256, 64), device="cuda").to(torch.bfloat16)
    dw = torch.zeros((27, 64), device="cuda"); db = torch.zeros(64, device="cuda")
    for _ in range(2):
        ops.conv_wgrad(ops.CONV_K3S1_C3, x, dy, dw, dbias=db)
torch.cuda.synchronize()
