"""CPU: the tiling / tap-table / weight-packing logic of the implicit-GEMM kernels (host code shared with the CUDA
kernels through b2g_conv_plan_query / b2g_pack_source_index) reproduces the oracle's Keras-semantics convolutions
when the GEMM itself is emulated with numpy."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import keras_ops as K

K3S1, K4S2, CONVT, K1S1 = 0, 1, 2, 3


def _plan(kind, n, h, w, tile=128):
    from imagegeneration_b200 import lib
    out = (C.c_int * 80)()
    rc = lib.load().b2g_conv_plan_query(kind, n, h, w, tile, out, 80)
    assert rc == 0, lib.load().b2g_last_error()
    v = list(out)
    keys = ["phases", "taps", "Hg", "Wg", "bw", "bh", "bn", "tiles_w", "tiles_h", "tiles_n", "m_tiles", "Hout", "Wout",
            "os", "n_amaps"]
    p = dict(zip(keys, v[:15]))
    p["tap_map"], p["tap_dy"], p["tap_dx"] = v[16:32], v[32:48], v[48:64]
    p["ph_y"], p["ph_x"] = v[64:68], v[68:72]
    return p


def _pack(kind, w, n_is_d3, flip, ck_pad=0):
    from imagegeneration_b200 import lib
    l = lib.load()
    kh, kw, d2, d3 = w.shape
    phases, taps = {K3S1: (1, 9), K4S2: (1, 16), CONVT: (4, 4), K1S1: (1, 1)}[kind]
    cn, ck = (d3, d2) if n_is_d3 else (d2, d3)
    ck = max(ck, ck_pad)
    flat = w.reshape(-1)
    out = np.zeros(phases * cn * taps * ck)
    for i in range(out.size):
        s = l.b2g_pack_source_index(kind, kh, kw, d2, d3, int(n_is_d3), int(flip), ck_pad, i)
        out[i] = 0.0 if s < 0 else flat[s]
    return out.reshape(phases, cn, taps * ck)


def _emulate(kind, x, packed, plan):
    """out[n, Y, X, :] accumulated tap by tap exactly as the kernel's TMA loads + MMAs do (zero fill outside)."""
    n, h, w, c = x.shape
    cn = packed.shape[1]
    out = np.zeros((n, plan["Hout"], plan["Wout"], cn))
    hg, wg = plan["Hg"], plan["Wg"]
    for ph in range(plan["phases"]):
        acc = np.zeros((n, hg, wg, cn))
        for t in range(plan["taps"]):
            e = ph * plan["taps"] + t
            m, dy, dx = plan["tap_map"][e], plan["tap_dy"][e], plan["tap_dx"][e]
            src = x[:, (m // 2)::2, (m % 2)::2, :] if plan["n_amaps"] == 4 else x
            sh = np.zeros((n, hg, wg, c))
            ys = np.arange(hg) + dy
            xs = np.arange(wg) + dx
            vy = (ys >= 0) & (ys < src.shape[1])
            vx = (xs >= 0) & (xs < src.shape[2])
            sh[np.ix_(np.arange(n), np.nonzero(vy)[0], np.nonzero(vx)[0])] = src[np.ix_(np.arange(n), ys[vy], xs[vx])]
            acc += sh @ packed[ph][:, t * c:(t + 1) * c].T
        out[:, plan["ph_y"][ph]::plan["os"], plan["ph_x"][ph]::plan["os"], :] = acc
    return out


def _rand(shape, seed):
    return np.random.default_rng(seed).uniform(-1, 1, shape)


@pytest.mark.parametrize("n,h,w", [(2, 4, 8), (1, 6, 4), (3, 2, 16)])
def test_k3s1_forward_and_dgrad(n, h, w):
    x, wt = _rand((n, h, w, 5), 1), _rand((3, 3, 5, 7), 2)
    ref = K.conv2d_same(torch.from_numpy(x), torch.from_numpy(wt), None, 1).numpy()
    plan = _plan(K3S1, n, h, w)
    assert np.allclose(_emulate(K3S1, x, _pack(K3S1, wt, True, False), plan), ref, atol=1e-10)
    dy = _rand((n, h, w, 7), 3)
    xt = torch.zeros((n, h, w, 5), dtype=torch.float64, requires_grad=True)
    K.conv2d_same(xt, torch.from_numpy(wt), None, 1).backward(torch.from_numpy(dy))
    assert np.allclose(_emulate(K3S1, dy, _pack(K3S1, wt, False, True), plan), xt.grad.numpy(), atol=1e-10)


@pytest.mark.parametrize("n,h,w", [(2, 4, 8), (1, 6, 4)])
def test_k4s2_forward_and_convt_dgrad(n, h, w):
    x, wt = _rand((n, 2 * h, 2 * w, 5), 4), _rand((4, 4, 5, 7), 5)
    ref = K.conv2d_same(torch.from_numpy(x), torch.from_numpy(wt), None, 2).numpy()
    assert np.allclose(_emulate(K4S2, x, _pack(K4S2, wt, True, False), _plan(K4S2, n, 2 * h, 2 * w)), ref, atol=1e-10)
    # input-gradient of the stride-2 conv = CONVT role with (n = Cin, k = Cout) packing
    dy = _rand((n, h, w, 7), 6)
    xt = torch.zeros((n, 2 * h, 2 * w, 5), dtype=torch.float64, requires_grad=True)
    K.conv2d_same(xt, torch.from_numpy(wt), None, 2).backward(torch.from_numpy(dy))
    got = _emulate(CONVT, dy, _pack(CONVT, wt, False, False), _plan(CONVT, n, h, w))
    assert np.allclose(got, xt.grad.numpy(), atol=1e-10)


@pytest.mark.parametrize("n,h,w", [(2, 4, 8), (1, 2, 4)])
def test_convt_forward_and_its_dgrad(n, h, w):
    x, wt = _rand((n, h, w, 5), 7), _rand((4, 4, 7, 5), 8)      # Conv2DTranspose kernel [kh,kw,Cout,Cin]
    xt = torch.from_numpy(x).requires_grad_(True)
    y = K.conv2d_transpose_same(xt, torch.from_numpy(wt), 2)
    assert np.allclose(_emulate(CONVT, x, _pack(CONVT, wt, False, False), _plan(CONVT, n, h, w)), y.detach().numpy(),
                       atol=1e-10)
    dy = _rand((n, 2 * h, 2 * w, 7), 9)
    y.backward(torch.from_numpy(dy))
    got = _emulate(K4S2, dy, _pack(K4S2, wt, True, False), _plan(K4S2, n, 2 * h, 2 * w))
    assert np.allclose(got, xt.grad.numpy(), atol=1e-10)


def test_k1s1_padding_of_the_edge_layer_weights():
    wt = _rand((3, 3, 3, 64), 10)
    p = _pack(K1S1, wt.reshape(1, 1, 27, 64), True, False, ck_pad=64)
    assert p.shape == (1, 64, 64)
    assert np.array_equal(p[0][:, :27], wt.reshape(27, 64).T) and not p[0][:, 27:].any()


@pytest.mark.parametrize("kind,h,w", [(K3S1, 144, 256), (K3S1, 72, 128), (K3S1, 36, 64), (K3S1, 18, 32), (K4S2, 144, 256),
                                      (CONVT, 18, 32), (CONVT, 72, 128), (K3S1, 2, 4), (K1S1, 144, 256)])
@pytest.mark.parametrize("n", [1, 3, 64])
@pytest.mark.parametrize("tile", [128, 64])
def test_tiles_cover_the_grid(kind, h, w, n, tile):
    p = _plan(kind, n, h, w, tile)
    assert p["bw"] * p["bh"] * p["bn"] == tile
    assert p["tiles_w"] * p["bw"] == p["Wg"]
    assert p["tiles_h"] * p["bh"] >= p["Hg"] and (p["tiles_h"] - 1) * p["bh"] < p["Hg"]
    assert p["tiles_n"] * p["bn"] >= n and (p["tiles_n"] - 1) * p["bn"] < n
    assert p["m_tiles"] == p["tiles_w"] * p["tiles_h"] * p["tiles_n"]
    assert max(p["bw"], p["bh"], p["bn"]) <= 256                      # TMA box limit


# ---------------------------------------------------------------------------------------------------------------
# row-slab schedule (csrc/b2g_conv_rows.cuh): slabs with a 1-pixel halo, dx taps as shifted views, R rows x P phases
# ---------------------------------------------------------------------------------------------------------------
def _rows_plan(kind, n, h, w, cin, block_n, pg):
    from imagegeneration_b200 import lib
    out = (C.c_int * 464)()
    rc = lib.load().b2g_conv_rows_plan_query(kind, n, h, w, cin, block_n, pg, out, 464)
    assert rc == 0, lib.load().b2g_last_error()
    v = list(out)
    d = dict(zip(["handled", "R", "P", "PG", "n_acc", "row_blocks", "n_slabs", "n_steps"], v[:8]))
    d["slabs"] = [tuple(v[8 + 3 * i:11 + 3 * i]) for i in range(d["n_slabs"])]
    d["steps"] = [tuple(v[80 + 6 * j:86 + 6 * j]) for j in range(d["n_steps"])]
    return d


def _emulate_rows(kind, x, packed, plan, block_n, ring):
    """Walks the schedule exactly like the kernel: slabs are 130-pixel rows (zero outside the image), MMA row m of
    tap dx reads slab pixel m + dxi, the slab ring is FIFO with `ring` stages."""
    n, h, w, cin = x.shape
    cn = packed.shape[1]
    assert cn <= block_n
    out = np.full((n, plan["Hout"], plan["Wout"], cn), np.nan)
    hg, wg = plan["Hg"], plan["Wg"]
    pg0 = _rows_plan(kind, n, h, w, cin, block_n, 0)
    assert pg0["handled"] == 1
    for pg in range(pg0["PG"]):
        rp = _rows_plan(kind, n, h, w, cin, block_n, pg)
        R, P = rp["R"], rp["P"]
        assert rp["row_blocks"] * R >= hg
        for img in range(n):
            for yb in range(rp["row_blocks"]):
                for tw in range(wg // 128):
                    y0, w0 = yb * R, tw * 128
                    slabs = []
                    for (m, dy, ch) in rp["slabs"]:
                        src = x[img, (m // 2)::2, (m % 2)::2, :] if plan["n_amaps"] == 4 else x[img]
                        sl = np.zeros((130, 64))
                        yy = y0 + dy
                        if 0 <= yy < src.shape[0]:
                            xs = np.arange(130) + w0 - 1
                            ok = (xs >= 0) & (xs < src.shape[1])
                            sl[ok] = src[yy, xs[ok], ch * 64:(ch + 1) * 64]
                        slabs.append(sl)
                    acc = np.zeros((R * P, 128, cn))
                    released = 0
                    for (bcol, phl, slab0, dxi, acc0, rel) in rp["steps"]:
                        bt = packed[pg * P + phl][:, bcol * 64:(bcol + 1) * 64]
                        for r in range(R):
                            s = slab0 + r
                            assert released <= s < released + ring          # still resident in the ring
                            acc[acc0 + r * P] += slabs[s][dxi:dxi + 128] @ bt.T
                        released += rel
                    assert released == len(slabs)
                    for a in range(R * P):
                        r, ph = a // P, pg * P + a % P
                        if y0 + r < hg:
                            oy = (y0 + r) * plan["os"] + plan["ph_y"][ph]
                            ox = (w0 + np.arange(128)) * plan["os"] + plan["ph_x"][ph]
                            out[img, oy, ox, :] = acc[a]
    return out


@pytest.mark.parametrize("kind,h,w,cin,cout,bn", [(K3S1, 5, 128, 128, 7, 128), (K3S1, 4, 256, 64, 3, 32),
                                                  (K3S1, 3, 128, 64, 9, 64), (K4S2, 6, 256, 64, 5, 128),
                                                  (CONVT, 3, 128, 128, 6, 64), (CONVT, 4, 128, 64, 5, 128)])
def test_row_slab_schedule_matches_the_tap_plan(kind, h, w, cin, cout, bn):
    n = 2
    ks = 3 if kind == K3S1 else 4
    x = _rand((n, h, w, cin), 11)
    if kind == CONVT:
        wt = _rand((ks, ks, cout, cin), 12)
        packed = _pack(CONVT, wt, False, False)
    else:
        wt = _rand((ks, ks, cin, cout), 12)
        packed = _pack(kind, wt, True, False)
    plan = _plan(kind, n, h, w)
    ref = _emulate(kind, x, packed, plan)
    got = _emulate_rows(kind, x, packed, plan, bn, ring=7 if bn == 128 else 8)
    assert not np.isnan(got).any()
    assert np.allclose(got, ref, atol=1e-9)


def test_row_slab_schedule_geometry_of_the_model_layers():
    # D.conv3 (k3s1 128->128 @72x128): 2 rows per super-tile, 4 slabs per channel chunk, 18 weight tiles
    p = _rows_plan(K3S1, 64, 72, 128, 128, 128, 0)
    assert (p["handled"], p["R"], p["P"], p["PG"], p["n_slabs"], p["n_steps"]) == (1, 2, 1, 1, 8, 18)
    # G.convT3 / D.conv2 dgrad (convT 128->64 @72x128): 2 rows x 2 phases, phase group = output row parity
    p = _rows_plan(CONVT, 64, 72, 128, 128, 64, 1)
    assert (p["handled"], p["R"], p["P"], p["PG"], p["n_slabs"], p["n_steps"]) == (1, 2, 2, 2, 6, 16)
    # G.convT4 (k3s1 64->3, N tile 32): 4 rows per super-tile share 6 slabs
    p = _rows_plan(K3S1, 64, 144, 256, 64, 32, 0)
    assert (p["handled"], p["R"], p["n_slabs"], p["n_steps"]) == (1, 4, 6, 9)
    # D.conv2 forward (k4s2 64->128): 4 parity maps x 3 slabs
    p = _rows_plan(K4S2, 64, 144, 256, 64, 128, 0)
    assert (p["handled"], p["R"], p["n_slabs"], p["n_steps"]) == (1, 2, 12, 16)
    # narrow layers stay on the per-tap kernel
    assert _rows_plan(K3S1, 64, 36, 64, 256, 128, 0)["handled"] == 0
    assert _rows_plan(K1S1, 64, 144, 256, 64, 64, 0)["handled"] == 0


def test_merged_rows_schedule_covers_the_same_taps():
    """RowsStep::wide (transposed conv, two sub-pixel phases per group): the merged schedule of the G.convT3 / D.conv2-dgrad
    role issues the dx = 0 shift of every dy ONCE for both phases (N = 2 x 64), first of its dy, and still covers every
    (phase, weight k-block) of the unmerged schedule exactly once; the slab ring of 6 divides the slabs, 3 weight stages
    divide the 12 steps (the compile-time kernel's static_asserts rely on both)."""
    from imagegeneration_b200 import lib
    kind, n, h, w, cin, bn = 2, 4, 72, 128, 128, 64      # B2G_CONVT_K4S2
    for pg in (0, 1):
        base = _rows_plan(kind, n, h, w, cin, bn, pg)
        out = (C.c_int * (8 + 72 + 8 * 64))()
        assert lib.load().b2g_conv_rows_plan_query_merged(kind, n, h, w, cin, bn, pg, 6, out, len(out)) == 0
        v = list(out)
        assert v[0] == 1 and v[1:6] == [base["R"], base["P"], base["PG"], base["n_acc"], base["row_blocks"]]
        assert v[6] == base["n_slabs"] == 6 and v[7] == 12 and base["n_steps"] == 16
        steps = [tuple(v[80 + 8 * j:88 + 8 * j]) for j in range(v[7])]
        want = sorted((phl, bcol, slab0, dxi) for (bcol, phl, slab0, dxi, acc0, rel) in base["steps"])
        got = []
        for (bcol, phl, slab0, dxi, acc0, rel, wide, bcol2) in steps:
            assert acc0 == phl
            got.append((phl, bcol, slab0, dxi))
            if wide:
                assert phl == 0 and dxi == 1                      # the shared shift is dx = 0
                got.append((1, bcol2, slab0, dxi))
        assert sorted(got) == want
        assert sum(s[6] for s in steps) == 4                      # 2 dy x 2 channel chunks
        assert steps[0][6] == 1                                   # a fresh super-tile starts with a wide step
        assert sum(s[5] for s in steps) == v[6]                   # every slab released exactly once
