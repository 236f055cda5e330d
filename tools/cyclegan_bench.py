"""CycleGAN train_step timing on one B200 (BASELINE configs[4]; the reference trains at 128 x 128, cyclegan/Trainer.py:5,26)
with the oracle's CPU step beside it.  python tools/cyclegan_bench.py [batch] [size] [steps]  -> one JSON line.
This path is the FIRST correct CUDA path of SURVEY row N2 (gather + GEMM), measured, not tuned."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200 import lib
from imagegeneration_b200.cyclegan import CycleGAN

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
S = int(sys.argv[2]) if len(sys.argv) > 2 else 128
K = int(sys.argv[3]) if len(sys.argv) > 3 else 5
size = (S, S, 3)
eng = CycleGAN(B, size)
g = torch.Generator().manual_seed(1)
x = (torch.rand((B, *size), generator=g) * 2 - 1).cuda()
y = (torch.rand((B, *size), generator=g) * 2 - 1).cuda()
l0 = lib.launch_count
out = eng.train_step(x, y)                 # eager: counts the launches of one step
per_step = lib.launch_count - l0
for _ in range(3):                         # the third step captures the CUDA graph, later ones replay it
    out = eng.train_step(x, y)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(K):
    out = eng.train_step(x, y)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / K
line = {"metric": "CycleGAN train images/sec (pairs)", "value": B / (ms * 1e-3), "unit": "image pairs/s", "n_gpus": 1,
        "steps": K, "ms_per_step": ms, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"CycleGAN train_step {S}x{S}, batch {B} (6 G + 6 D forwards, 4 Adam updates)"},
        "gpu_launches": per_step * K, "graph_replay": eng._graph is not None, "losses": {k: round(float(v), 5) for k, v in out.items()},
        "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9}
if "--no-cpu" not in sys.argv:
    from oracle import cyclegan_oracle as CO
    torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    orc = CO.OracleCycleGAN(size)
    xb, yb = x[:1].cpu(), y[:1].cpu()
    orc.train_step(xb, yb)
    t0 = time.perf_counter()
    orc.train_step(xb, yb)
    dt = time.perf_counter() - t0
    line["cpu_baseline"] = {"value": 1 / dt, "unit": "image pairs/s", "cores": torch.get_num_threads(), "kind": "port",
                            "sample": f"1 oracle train_step at batch 1, {S}x{S}, {dt:.2f} s"}
print(json.dumps(line))
