"""A handful of representative conv_igemm / wgrad launches at the bench shapes (B=64), for `ncu --set full`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imagegeneration_b200 import ops

B = 64
dev = "cuda"


def rnd(*shape, dtype=torch.bfloat16):
    return (torch.rand(shape, device=dev) * 2 - 1).to(dtype)


reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
for _ in range(reps):
    # 0. D.conv1 forward with the im2col patch built in shared memory (B2G_CONV_K3S1_C3)
    img = rnd(B, 144, 256, 3, dtype=torch.float32)
    wc3 = ops.pack_conv_weight(ops.CONV_K1S1, rnd(1, 1, 27, 64, dtype=torch.float32), n_is_d3=True, ck_pad=64)
    ops.conv_igemm(ops.CONV_K3S1_C3, img, wc3, bias=rnd(64, dtype=torch.float32), act=ops.ACT_LEAKY, alpha=0.1,
                   dropout_thresh8=128, dropout_seed=4)
    # 1. D.conv1 forward as K1S1 GEMM (im2col27 operand): bias + leaky + dropout
    x27 = rnd(B, 144, 256, 64)
    w = ops.pack_conv_weight(ops.CONV_K1S1, rnd(1, 1, 27, 64, dtype=torch.float32), n_is_d3=True, ck_pad=64)
    bias = rnd(64, dtype=torch.float32)
    ops.conv_igemm(ops.CONV_K1S1, x27, w, bias=bias, act=ops.ACT_LEAKY, alpha=0.1, dropout_thresh8=128, dropout_seed=5)
    # 2. D.conv2 dgrad = convT 128 -> 64 with LeakyReLU/dropout derivative epilogue
    dy = rnd(B, 72, 128, 128)
    w2 = ops.pack_conv_weight(ops.CONVT_K4S2, rnd(4, 4, 64, 128, dtype=torch.float32), n_is_d3=False)
    yprev = rnd(B, 144, 256, 64)
    ops.conv_igemm(ops.CONVT_K4S2, dy, w2, yprev=yprev, dact=ops.ACT_LEAKY, dact_alpha=0.1, dact_scale=2.0)
    # 3. D.conv3 forward 128 -> 128 k3s1
    x3 = rnd(B, 72, 128, 128)
    w3 = ops.pack_conv_weight(ops.CONV_K3S1, rnd(3, 3, 128, 128, dtype=torch.float32), n_is_d3=True)
    b3 = rnd(128, dtype=torch.float32)
    ops.conv_igemm(ops.CONV_K3S1, x3, w3, bias=b3, act=ops.ACT_LEAKY, alpha=0.1, dropout_thresh8=128, dropout_seed=6)
    # 4. D.conv7 forward 512 -> 512 k3s1
    x7 = rnd(B, 18, 32, 512)
    w7 = ops.pack_conv_weight(ops.CONV_K3S1, rnd(3, 3, 512, 512, dtype=torch.float32), n_is_d3=True)
    b7 = rnd(512, dtype=torch.float32)
    ops.conv_igemm(ops.CONV_K3S1, x7, w7, bias=b7, act=ops.ACT_LEAKY, alpha=0.1, dropout_thresh8=128, dropout_seed=7)
    # 5. G.convT4 forward 64 -> 3 + tanh (narrow fp32 output)
    a3 = rnd(B, 144, 256, 64)
    w4 = ops.pack_conv_weight(ops.CONV_K3S1, rnd(3, 3, 3, 64, dtype=torch.float32), n_is_d3=False, flip=True)
    ops.conv_igemm(ops.CONV_K3S1, a3, w4, act=ops.ACT_TANH, out_f32=True)
    # 6. wgrad conv3 (128/128) and conv5 (256/256)
    dw3 = torch.zeros(3, 3, 128, 128, device=dev)
    ops.conv_wgrad(ops.CONV_K3S1, x3, rnd(B, 72, 128, 128), dw3)
    dw5 = torch.zeros(3, 3, 256, 256, device=dev)
    ops.conv_wgrad(ops.CONV_K3S1, rnd(B, 36, 64, 256), rnd(B, 36, 64, 256), dw5)
    # 7. wgrad conv2 (64 -> 128, stride 2): tap-grouped kernel, 16 taps
    dw2 = torch.zeros(4, 4, 64, 128, device=dev)
    ops.conv_wgrad(ops.CONV_K4S2, rnd(B, 144, 256, 64), rnd(B, 72, 128, 128), dw2)
torch.cuda.synchronize()
print("ok")
