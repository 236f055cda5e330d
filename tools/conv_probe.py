"""Times the narrow (Cout <= 128) conv roles of the step at the bench shape, normally and with the epilogue switched off
(B2G_DBG_EPI=1: accumulators are drained, no math / saved-output loads / stores) — the second number is the MMA + operand
pipeline alone, so the difference says how much of a launch the epilogue costs.  Run on the GPU box."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def rnd(*shape):
    return (torch.rand(shape, device="cuda") * 2 - 1).to(torch.bfloat16)


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def both(name, fn, flops):
    os.environ["B2G_DBG_EPI"] = "0"
    t = timeit(fn)
    os.environ["B2G_DBG_EPI"] = "1"
    t0 = timeit(fn)
    os.environ["B2G_DBG_EPI"] = "0"
    print(f"{name:46s} {t:7.1f} us {flops / t / 1e6:7.1f} TF/s | no-epilogue {t0:7.1f} us {flops / t0 / 1e6:7.1f} TF/s", flush=True)


W = lambda *s: torch.rand(*s, device="cuda") - 0.5
x128 = rnd(B, 72, 128, 128); x64 = rnd(B, 144, 256, 64); x256 = rnd(B, 36, 64, 256)
o64 = torch.empty((B, 144, 256, 64), device="cuda", dtype=torch.bfloat16); y64 = rnd(B, 144, 256, 64)
o128 = torch.empty((B, 72, 128, 128), device="cuda", dtype=torch.bfloat16); y128 = rnd(B, 72, 128, 128)
b64 = torch.zeros(64, device="cuda"); b128 = torch.zeros(128, device="cuda")
cs64 = torch.zeros(64, device="cuda"); cs128 = torch.zeros(128, device="cuda")
fl = 2.0 * B * 144 * 256 * 64 * 128 * 4
wt3 = ops.pack_conv_weight(ops.CONVT_K4S2, W(4, 4, 64, 128), n_is_d3=False)
both("convT3 fwd 128->64 (bias, relu)", lambda: ops.conv_igemm(ops.CONVT_K4S2, x128, wt3, o64, bias=b64, act=ops.ACT_RELU), fl)
both("D.conv2 dgrad 128->64 (leaky', colsum)", lambda: ops.conv_igemm(ops.CONVT_K4S2, x128, wt3, o64, yprev=y64, dact=ops.ACT_LEAKY,
                                                                      dact_alpha=0.1, dact_scale=2.0, colsum=cs64), fl)
wc2 = ops.pack_conv_weight(ops.CONV_K4S2, W(4, 4, 64, 128), n_is_d3=True)
both("D.conv2 fwd 64->128 s2 (bias, leaky, dropout)", lambda: ops.conv_igemm(ops.CONV_K4S2, x64, wc2, o128, bias=b128, act=ops.ACT_LEAKY,
                                                                             alpha=0.1, dropout_thresh8=128, dropout_seed=3), fl)
both("convT3 dgrad 64->128 s2 (relu', colsum)", lambda: ops.conv_igemm(ops.CONV_K4S2, x64, wc2, o128, yprev=y128, dact=ops.ACT_RELU,
                                                                       colsum=cs128), fl)
w3 = ops.pack_conv_weight(ops.CONV_K3S1, W(3, 3, 128, 128), n_is_d3=True)
fl3 = 2.0 * B * 72 * 128 * 128 * 128 * 9
both("D.conv3 fwd 128->128 (bias, leaky, dropout)", lambda: ops.conv_igemm(ops.CONV_K3S1, x128, w3, o128, bias=b128, act=ops.ACT_LEAKY,
                                                                           alpha=0.1, dropout_thresh8=128, dropout_seed=3), fl3)
both("D.conv3 dgrad (leaky', colsum)", lambda: ops.conv_igemm(ops.CONV_K3S1, x128, w3, o128, yprev=y128, dact=ops.ACT_LEAKY,
                                                              dact_alpha=0.1, dact_scale=2.0, colsum=cs128), fl3)
wt2 = ops.pack_conv_weight(ops.CONVT_K4S2, W(4, 4, 128, 256), n_is_d3=False)
fl2 = 2.0 * B * 72 * 128 * 128 * 256 * 4
both("convT2 fwd 256->128 (bias, relu)", lambda: ops.conv_igemm(ops.CONVT_K4S2, x256, wt2, o128, bias=b128, act=ops.ACT_RELU), fl2)
both("D.conv4 dgrad 256->128 (leaky', colsum)", lambda: ops.conv_igemm(ops.CONVT_K4S2, x256, wt2, o128, yprev=y128, dact=ops.ACT_LEAKY,
                                                                       dact_alpha=0.1, dact_scale=2.0, colsum=cs128), fl2)
x3 = torch.rand((B, 144, 256, 3), device="cuda") * 2 - 1
wc1 = ops.pack_conv_weight(ops.CONV_K1S1, W(1, 1, 27, 64), n_is_d3=True, ck_pad=64)
flc = 2.0 * B * 144 * 256 * 27 * 64
both("conv_c3 fwd 3->64 (bias, leaky, dropout)", lambda: ops.conv_igemm(ops.CONV_K3S1_C3, x3, wc1, o64, bias=b64, act=ops.ACT_LEAKY, alpha=0.1,
                                                                        dropout_thresh8=128, dropout_seed=7), flc)
both("conv_c3 dgrad role (relu', colsum)", lambda: ops.conv_igemm(ops.CONV_K3S1_C3, x3, wc1, o64, yprev=y64, dact=ops.ACT_RELU, colsum=cs64), flc)
