"""Runs the 3-channel edge kernels of the step a few times at the bench shape (for ncu captures / quick timings):
   python tools/edge_probe.py [B]   -> conv_c3 fwd (bias, leaky, dropout), conv_c3 dgrad role, conv_o3 fwd + dgrad role,
   wgrad_c3."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200 import ops
from imagegeneration_b200.sndcgan import SNDCGANEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
eng = SNDCGANEngine(B, dropout=0.5)
D, G = eng.disc_model, eng.gen_model
D.repack(); G.repack()
x = torch.rand((B, 144, 256, 3), device="cuda") * 2 - 1
y = torch.empty((B, 144, 256, 64), device="cuda", dtype=torch.bfloat16)
yprev = (torch.rand((B, 144, 256, 64), device="cuda") - 0.3).to(torch.bfloat16)
bias = torch.zeros(64, device="cuda")
g3 = torch.empty((B, 144, 256, 3), device="cuda")
dw = torch.zeros((27, 64), device="cuda")
db = torch.zeros(64, device="cuda")
cs = torch.zeros(64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
px = B * 144 * 256


def run(name, fn, byt):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    us = ts[len(ts) // 2]
    print(f"{name:40s} {us:8.1f} us  {byt / us / 1e3:7.1f} GB/s", flush=True)


run("conv_c3 fwd (bias, leaky, dropout)", lambda: ops.conv_igemm(ops.CONV_K3S1_C3, x, D._wp_f[0], y, bias=bias, act=ops.ACT_LEAKY,
                                                                 alpha=0.1, dropout_thresh8=128, dropout_seed=7), px * 140)
run("conv_c3 dgrad role (relu', colsum)", lambda: ops.conv_igemm(ops.CONV_K3S1_C3, x, G._wp4_d, y, yprev=yprev, dact=ops.ACT_RELU,
                                                                 colsum=cs), px * 268)
run("conv_o3 fwd 64->3 (tanh)", lambda: ops.conv_igemm(ops.CONV_K3S1, yprev, G._wp4_f, g3, act=ops.ACT_TANH, out_f32=True), px * 140)
run("conv_o3 dgrad role (tanh')", lambda: ops.conv_igemm(ops.CONV_K3S1, yprev, D._wp_d[0], g3, yprev=x, dact=ops.ACT_TANH,
                                                         out_f32=True), px * 152)
run("wgrad_c3 (+bias)", lambda: ops.conv_wgrad(ops.CONV_K3S1_C3, x, yprev, dw, dbias=db), px * 140)
