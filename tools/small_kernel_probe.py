"""CUDA-event timings of the small HBM-bound kernels of the step at the bench shape (run on the GPU box):
discriminator head backward, Adam, weight packing.  python tools/small_kernel_probe.py"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200 import lib as L
from imagegeneration_b200.sndcgan import SNDCGANEngine

B = 64
lib = L.load()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
p = lambda t: C.c_void_p(t.data_ptr())
st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)


def timeit(name, fn, bytes_moved, reps=10):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.sum()                       # read flush: no dirty lines draining into the timed kernel
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    us = ts[len(ts) // 2]
    print(f"{name:44s} {us:8.1f} us  {bytes_moved / us / 1e3:7.1f} GB/s", flush=True)


F = 18 * 32 * 512
x = (torch.rand((B, F), device="cuda") - 0.5).to(torch.bfloat16)
w = torch.rand(F, device="cuda") - 0.5
dl = torch.rand(B, device="cuda") - 0.5
dx = torch.empty_like(x)
dw = torch.zeros(F, device="cuda")
db = torch.zeros(1, device="cuda")
timeit("dhead_bwd (B=64, F=294912, bf16)",
       lambda: L.check(lib.b2g_dhead_bwd(p(x), p(w), None, p(dl), B, F, 0.1, 2.0, 0.0, p(dx), p(dw), p(db), st()), "dhead_bwd"),
       2 * B * F * 2 + 8 * F)
eng = SNDCGANEngine(B, dropout=0.5, cuda_graph=False)
D, G = eng.disc_model, eng.gen_model
timeit("pack all discriminator weights (14 packings)", lambda: D.repack(), D.arena.numel * 4 + 2 * D.arena.numel * 2)
timeit("pack all generator weights (8 packings)", lambda: G.repack(), 3.4e6 * 4 + 2 * 3.4e6 * 2)
