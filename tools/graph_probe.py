"""Measurement only: how much of a step is launch gaps?  Replays one captured train_step (same dropout seeds every
replay, so NOT a valid training loop) and compares with eager launches."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imagegeneration_b200.sndcgan import SNDCGANEngine

B = 64
eng = SNDCGANEngine(B, dropout=0.5)
real = torch.rand((B, 144, 256, 3), device="cuda") * 2 - 1
z = torch.rand((B, 128), device="cuda") * 2 - 1
for _ in range(3):
    eng.train_step(real, z)
torch.cuda.synchronize()

def timed(fn, n=20):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    a.record()
    for _ in range(n):
        fn()
    t1 = time.perf_counter()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n, (t1 - t0) * 1e3 / n

ms, host = timed(lambda: eng.train_step(real, z))
print(f"eager: {ms:.3f} ms/step on the GPU, host enqueue {host:.3f} ms/step")
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    eng.train_step(real, z)
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=s):
        eng.train_step(real, z)
torch.cuda.current_stream().wait_stream(s)
ms, host = timed(g.replay)
print(f"graph: {ms:.3f} ms/step on the GPU, host enqueue {host:.3f} ms/step")
