"""One launch each of the epilogue-bound roles, for `ncu --set full --import-source on` (run on the GPU box):
D.conv2 dgrad (conv_rows<64>, leaky' + colsum), conv_c3 forward (bias, leaky, dropout), conv_c3 dgrad role (relu', colsum),
D.conv4 dgrad (conv_pair<128>)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200 import ops

B = 64
rnd = lambda *s: (torch.rand(s, device="cuda") * 2 - 1).to(torch.bfloat16)
W = lambda *s: torch.rand(*s, device="cuda") - 0.5
x128 = rnd(B, 72, 128, 128); o64 = torch.empty((B, 144, 256, 64), device="cuda", dtype=torch.bfloat16); y64 = rnd(B, 144, 256, 64)
cs64 = torch.zeros(64, device="cuda"); b64 = torch.zeros(64, device="cuda")
wt3 = ops.pack_conv_weight(ops.CONVT_K4S2, W(4, 4, 64, 128), n_is_d3=False)
ops.conv_igemm(ops.CONVT_K4S2, x128, wt3, o64, yprev=y64, dact=ops.ACT_LEAKY, dact_alpha=0.1, dact_scale=2.0, colsum=cs64)
x3 = torch.rand((B, 144, 256, 3), device="cuda") * 2 - 1
wc1 = ops.pack_conv_weight(ops.CONV_K1S1, W(1, 1, 27, 64), n_is_d3=True, ck_pad=64)
ops.conv_igemm(ops.CONV_K3S1_C3, x3, wc1, o64, bias=b64, act=ops.ACT_LEAKY, alpha=0.1, dropout_thresh8=128, dropout_seed=7)
ops.conv_igemm(ops.CONV_K3S1_C3, x3, wc1, o64, yprev=y64, dact=ops.ACT_RELU, colsum=cs64)
x256 = rnd(B, 36, 64, 256); o128 = torch.empty((B, 72, 128, 128), device="cuda", dtype=torch.bfloat16); y128 = rnd(B, 72, 128, 128)
cs128 = torch.zeros(128, device="cuda")
wt2 = ops.pack_conv_weight(ops.CONVT_K4S2, W(4, 4, 128, 256), n_is_d3=False)
ops.conv_igemm(ops.CONVT_K4S2, x256, wt2, o128, yprev=y128, dact=ops.ACT_LEAKY, dact_alpha=0.1, dact_scale=2.0, colsum=cs128)
torch.cuda.synchronize()
print("done")
# conv_pair<256> roles: D.conv6 input gradient (transposed, 512 -> 256, 4 phases) and D.conv5 forward (256 -> 256)
x512 = rnd(B, 18, 32, 512); o256 = torch.empty((B, 36, 64, 256), device="cuda", dtype=torch.bfloat16); y256 = rnd(B, 36, 64, 256)
cs256 = torch.zeros(256, device="cuda"); b256 = torch.zeros(256, device="cuda")
wt6 = ops.pack_conv_weight(ops.CONVT_K4S2, W(4, 4, 256, 512), n_is_d3=False)
ops.conv_igemm(ops.CONVT_K4S2, x512, wt6, o256, yprev=y256, dact=ops.ACT_LEAKY, dact_alpha=0.1, dact_scale=2.0, colsum=cs256)
w5 = ops.pack_conv_weight(ops.CONV_K3S1, W(3, 3, 256, 256), n_is_d3=True)
ops.conv_igemm(ops.CONV_K3S1, y256, w5, o256, bias=b256, act=ops.ACT_LEAKY, alpha=0.1, dropout_thresh8=128, dropout_seed=7)
torch.cuda.synchronize()
print("done 256")
