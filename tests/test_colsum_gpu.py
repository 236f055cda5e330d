"""GPU: per-channel column sums fused into backward conv epilogues (b2g_conv_args.colsum) — the bias / BatchNorm-beta
gradients of SNDCGAN.py:248,:256,:266 (tape.gradient) without a separate reduction pass.  Checked against the sum of
the tensor the same call stored, for every kernel family that can take the call (per-tap, CTA pair with and without
the TMA epilogue, row-slab static schedules, 3-channel edge kernel), plus ragged shapes with masked pixels."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(shape, generator=g) * 2 - 1) * scale


def _yprev(shape, seed):
    y = _rand(shape, seed)
    y[torch.rand(shape, generator=torch.Generator().manual_seed(seed + 1)) < 0.4] = 0.0
    return y.to(torch.bfloat16).cuda()


CASES = [
    # (kind name, N, H, W, Cin, Cout)   H, W = input grid of the call
    ("K3S1", 2, 18, 32, 512, 512),      # CTA pair, N = 256 tiles, direct stores
    ("K3S1", 64, 72, 128, 128, 128),    # row-slab static schedule (bench shape), TMA epilogue
    ("K3S1", 2, 7, 128, 128, 64),       # row-slab generic schedule, odd height
    ("K3S1", 3, 5, 16, 64, 64),         # per-tap kernel, ragged tiles (masked pixels)
    ("CONVT", 64, 72, 128, 128, 64),    # row-slab static convT (bench shape), 4 output phases
    ("CONVT", 2, 36, 64, 256, 128),     # CTA pair N = 128, TMA epilogue, 4 phases
    ("CONVT", 1, 9, 16, 64, 64),        # per-tap / small
    ("K4S2", 4, 144, 256, 64, 128),     # row-slab static stride-2 conv (direct stores)
    ("K4S2", 2, 36, 64, 256, 512),      # CTA pair stride-2
]


@pytest.mark.parametrize("kind,n,h,w,cin,cout", CASES)
@pytest.mark.parametrize("tma", ["1", "0"])
def test_fused_colsum_matches_stored_output(kind, n, h, w, cin, cout, tma):
    from imagegeneration_b200 import ops
    k = {"K3S1": ops.CONV_K3S1, "CONVT": ops.CONVT_K4S2, "K4S2": ops.CONV_K4S2}[kind]
    ksz = 3 if kind == "K3S1" else 4
    x = _rand((n, h, w, cin), 1).to(torch.bfloat16).cuda()
    if kind == "CONVT":
        wt = _rand((ksz, ksz, cout, cin), 2, 0.05).cuda()
        wp = ops.pack_conv_weight(k, wt, n_is_d3=False)
        oshape = (n, 2 * h, 2 * w, cout)
    else:
        wt = _rand((ksz, ksz, cin, cout), 2, 0.05).cuda()
        wp = ops.pack_conv_weight(k, wt, n_is_d3=True)
        oshape = (n, h // 2, w // 2, cout) if kind == "K4S2" else (n, h, w, cout)
    yp = _yprev(oshape, 3)
    cs = torch.full((cout,), 0.5, device="cuda")          # accumulated into: starts non-zero
    os.environ["B2G_EPI_TMA"] = tma
    try:
        out = ops.conv_igemm(k, x, wp, yprev=yp, dact=ops.ACT_LEAKY, dact_alpha=0.1, dact_scale=2.0, dact_zero=0.0,
                             colsum=cs)
        ref_out = ops.conv_igemm(k, x, wp, yprev=yp, dact=ops.ACT_LEAKY, dact_alpha=0.1, dact_scale=2.0, dact_zero=0.0)
    finally:
        os.environ.pop("B2G_EPI_TMA", None)
    torch.cuda.synchronize()
    assert torch.equal(out, ref_out)                      # the fused sums do not disturb the stored tensor
    want = out.double().sum(dim=(0, 1, 2)) + 0.5
    err = (cs.double() - want).abs().max() / (want.abs().max() + 1e-30)
    assert float(err) < 2e-5, float(err)


def test_fused_colsum_edge_kernel():
    """3-channel edge kernel (G.convT4 input gradient): sums of the stored dq = beta gradient of bn3."""
    from imagegeneration_b200 import ops
    n, h, w = 3, 16, 32
    x = _rand((n, h, w, 3), 5).cuda()
    wt = _rand((1, 1, 27, 64), 6, 0.1).cuda()
    wp = ops.pack_conv_weight(ops.CONV_K1S1, wt, n_is_d3=True, ck_pad=64)
    yp = _yprev((n, h, w, 64), 7)
    cs = torch.zeros(64, device="cuda")
    out = ops.conv_igemm(ops.CONV_K3S1_C3, x, wp, yprev=yp, dact=ops.ACT_RELU, colsum=cs)
    torch.cuda.synchronize()
    want = out.double().sum(dim=(0, 1, 2))
    assert float((cs.double() - want).abs().max() / want.abs().max()) < 2e-5


def test_colsum_rejected_on_forward_epilogues():
    from imagegeneration_b200 import ops
    from imagegeneration_b200.lib import B200GanError
    x = torch.zeros((1, 8, 16, 64), dtype=torch.bfloat16, device="cuda")
    wp = ops.pack_conv_weight(ops.CONV_K3S1, torch.zeros((3, 3, 64, 64), device="cuda"), n_is_d3=True)
    with pytest.raises(B200GanError):
        ops.conv_igemm(ops.CONV_K3S1, x, wp, bias=torch.zeros(64, device="cuda"), act=ops.ACT_LEAKY, alpha=0.1,
                       colsum=torch.zeros(64, device="cuda"))


@pytest.mark.parametrize("kind,n,h,w,cin,cout", [
    ("CONVT", 3, 6, 8, 512, 256),     # CTA pair N = 256, boxes for the backward role, ragged tiles (3 images, 6 rows)
    ("K3S1", 3, 12, 16, 256, 256),    # same, 3x3
    ("CONVT", 3, 10, 16, 256, 128),   # CTA pair N = 128, ragged
    ("K3S1", 3, 5, 128, 128, 128),    # row-slab kernel, odd height
])
def test_fused_colsum_ragged_tiles_with_nonzero_derivative_at_zero(kind, n, h, w, cin, cout):
    """act'(0) != 0 (layers without dropout, WGAN: dact_zero = alpha * scale): the pixels of a ragged tile that lie outside
    the tensor have non-zero accumulators (their taps reach into the image) and a zero-filled saved output does NOT mask
    them — the kernels must clear them before the box column sums (regression: 2.5-4 % error in round 2)."""
    from imagegeneration_b200 import ops
    k = {"K3S1": ops.CONV_K3S1, "CONVT": ops.CONVT_K4S2}[kind]
    x = _rand((n, h, w, cin), 1).to(torch.bfloat16).cuda()
    if kind == "CONVT":
        wp = ops.pack_conv_weight(k, _rand((4, 4, cout, cin), 2, 0.05).cuda(), n_is_d3=False)
        oshape = (n, 2 * h, 2 * w, cout)
    else:
        wp = ops.pack_conv_weight(k, _rand((3, 3, cin, cout), 2, 0.05).cuda(), n_is_d3=True)
        oshape = (n, h, w, cout)
    yp = _yprev(oshape, 3)
    cs = torch.zeros(cout, device="cuda")
    out = ops.conv_igemm(k, x, wp, yprev=yp, dact=ops.ACT_LEAKY, dact_alpha=0.1, dact_scale=1.0, dact_zero=0.1, colsum=cs)
    torch.cuda.synchronize()
    want = out.double().sum(dim=(0, 1, 2))
    err = (cs.double() - want).abs().max() / (want.abs().max() + 1e-30)
    assert float(err) < 2e-5, float(err)
