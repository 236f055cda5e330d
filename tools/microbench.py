"""CUDA-event timings of the memory-bound edge kernels at the bench shapes (run on the GPU box)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200 import ops
from imagegeneration_b200.sndcgan import SNDCGANEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(name, fn, bytes_moved=0, reps=10):
    fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    us = tot / reps * 1e3
    extra = f"  {bytes_moved / us / 1e3:7.1f} GB/s" if bytes_moved else ""
    print(f"{name:44s} {us:8.1f} us{extra}")


eng = SNDCGANEngine(B, dropout=0.5)
D, G = eng.disc_model, eng.gen_model
D.repack(); G.repack()
timeit("D.repack (14 packings, one launch)", D.repack)
timeit("G.repack (8 packings, one launch)", G.repack)

x = torch.rand((B, 144, 256, 3), device="cuda") * 2 - 1
y = torch.empty((B, 144, 256, 64), device="cuda", dtype=torch.bfloat16)
bias = torch.zeros(64, device="cuda")
px = B * 144 * 256
timeit("conv_c3 fwd (bias, leaky, dropout)",
       lambda: ops.conv_igemm(ops.CONV_K3S1_C3, x, D._wp_f[0], y, bias=bias, act=ops.ACT_LEAKY, alpha=0.1,
                              dropout_thresh8=128, dropout_seed=7), px * (12 + 128))
yprev = torch.rand((B, 144, 256, 64), device="cuda").to(torch.bfloat16)
timeit("conv_c3 dgrad role (relu' from yprev)",
       lambda: ops.conv_igemm(ops.CONV_K3S1_C3, x, G._wp4_d, y, yprev=yprev, dact=ops.ACT_RELU), px * (12 + 256))
x27 = torch.empty((B, 144, 256, 64), device="cuda", dtype=torch.bfloat16)
timeit("im2col27", lambda: ops.im2col27(x, x27), px * (12 + 128))
g3 = torch.empty((B, 144, 256, 3), device="cuda")
timeit("convT4 fwd 64->3 (tanh, fp32 out)",
       lambda: ops.conv_igemm(ops.CONV_K3S1, yprev, G._wp4_f, g3, act=ops.ACT_TANH, out_f32=True), px * (128 + 12))
dw = torch.zeros((27, 64), device="cuda")
db = torch.zeros(64, device="cuda")
timeit("edge wgrad (x27 x dy -> [27][64] + bias)",
       lambda: ops.conv_wgrad(ops.CONV_K1S1, x27, yprev, dw, ca_valid=27, dbias=db), px * 256)

# --- narrow tensor-core layers (epilogue-bound candidates)
def rnd(*shape):
    return (torch.rand(shape, device="cuda") * 2 - 1).to(torch.bfloat16)


def tf(name, fn, flops):
    fn(); torch.cuda.synchronize()
    tot = 0.0
    for _ in range(10):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    us = tot / 10 * 1e3
    print(f"{name:44s} {us:8.1f} us  {flops / us / 1e6:7.1f} TF/s")


f32 = torch.float32
x = rnd(B, 72, 128, 128); o = torch.empty((B, 144, 256, 64), device="cuda", dtype=torch.bfloat16)
w = ops.pack_conv_weight(ops.CONVT_K4S2, torch.rand(4, 4, 64, 128, device="cuda") - 0.5, n_is_d3=False)
bias = torch.zeros(64, device="cuda")
fl = 2.0 * B * 144 * 256 * 64 * 128 * 4
tf("convT3 fwd 128->64 (relu)", lambda: ops.conv_igemm(ops.CONVT_K4S2, x, w, o, bias=bias, act=ops.ACT_RELU), fl)
tf("convT3 role dgrad (leaky' from yprev)", lambda: ops.conv_igemm(ops.CONVT_K4S2, x, w, o, yprev=yprev, dact=ops.ACT_LEAKY,
                                                                    dact_alpha=0.1, dact_scale=2.0), fl)
x3 = rnd(B, 72, 128, 128); o3 = torch.empty_like(x3); y3 = rnd(B, 72, 128, 128)
w3 = ops.pack_conv_weight(ops.CONV_K3S1, torch.rand(3, 3, 128, 128, device="cuda") - 0.5, n_is_d3=True)
b3 = torch.zeros(128, device="cuda")
fl3 = 2.0 * B * 72 * 128 * 128 * 128 * 9
tf("conv3 fwd 128->128 (leaky, dropout)", lambda: ops.conv_igemm(ops.CONV_K3S1, x3, w3, o3, bias=b3, act=ops.ACT_LEAKY, alpha=0.1,
                                                                  dropout_thresh8=128, dropout_seed=3), fl3)
tf("conv3 dgrad (leaky' from yprev)", lambda: ops.conv_igemm(ops.CONV_K3S1, x3, w3, o3, yprev=y3, dact=ops.ACT_LEAKY,
                                                             dact_alpha=0.1, dact_scale=2.0), fl3)
x2 = rnd(B, 36, 64, 256)
w2 = ops.pack_conv_weight(ops.CONVT_K4S2, torch.rand(4, 4, 128, 256, device="cuda") - 0.5, n_is_d3=False)
fl2 = 2.0 * B * 72 * 128 * 128 * 256 * 4
tf("convT2 fwd 256->128 (relu)", lambda: ops.conv_igemm(ops.CONVT_K4S2, x2, w2, o3, bias=b3, act=ops.ACT_RELU), fl2)
tf("convT2 role dgrad (leaky' from yprev)", lambda: ops.conv_igemm(ops.CONVT_K4S2, x2, w2, o3, yprev=y3, dact=ops.ACT_LEAKY,
                                                                    dact_alpha=0.1, dact_scale=2.0), fl2)
xc2 = rnd(B, 144, 256, 64)
wc2 = ops.pack_conv_weight(ops.CONV_K4S2, torch.rand(4, 4, 64, 128, device="cuda") - 0.5, n_is_d3=True)
flc2 = 2.0 * B * 72 * 128 * 128 * 64 * 16
tf("conv2 fwd 64->128 s2 (leaky, dropout)", lambda: ops.conv_igemm(ops.CONV_K4S2, xc2, wc2, o3, bias=b3, act=ops.ACT_LEAKY, alpha=0.1,
                                                                    dropout_thresh8=128, dropout_seed=3), flc2)

# --- generator stem
import math
from imagegeneration_b200 import lib as L
F = 512 * 18 * 32
zz = torch.rand((B, 128), device="cuda") * 2 - 1
Wd = (torch.rand((128, F), device="cuda") - 0.5) * 0.1
gm = torch.ones(F, device="cuda"); bt = torch.zeros(F, device="cuda")
od = torch.empty((B, F), device="cuda", dtype=torch.bfloat16)
lib = L.load()
_p = lambda t: t.data_ptr()
_s = lambda: torch.cuda.current_stream().cuda_stream
timeit("stem fwd (W 151 MB + out)", lambda: L.check(lib.b2g_dense_g_fwd(_p(zz), _p(Wd), _p(gm), _p(bt), 0.9995, _p(od), B, F, _s()), "f"),
       F * 128 * 4 + B * F * 2)
dWd = torch.empty_like(Wd); dg = torch.empty(F, device="cuda"); dbt = torch.empty(F, device="cuda")
timeit("stem bwd (W + dW + dq)", lambda: L.check(lib.b2g_dense_g_bwd(_p(zz), _p(Wd), _p(gm), 0.9995, _p(od), _p(dWd), _p(dg), _p(dbt), B, F, _s()), "b"),
       F * 128 * 8 + B * F * 2)
x_img = torch.rand((B, 144, 256, 3), device="cuda") * 2 - 1
timeit("edge wgrad fused (fp32 x, dy -> [27][64] + bias)",
       lambda: ops.conv_wgrad(ops.CONV_K3S1_C3, x_img, yprev, dw, dbias=db), px * (12 + 128))
