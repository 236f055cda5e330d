"""b2g_dhead_fwd_loss_static — the allocation-free variant of the discriminator head (Flatten + Dense(1) + loss,
sndcgan/SNDCGAN.py:122-123,:170) whose partial-sum scratch is a static device array inside libb200gan.so.  It has no
counterpart in the reference; it exists for callers that cannot call cudaMalloc (include/b200gan.h).

Called through the C ABI and compared with the existing entry point b2g_dhead_fwd_loss (caller-supplied workspace) on
two calls in a row with different batches / inputs — the second call sees the first call's leftovers in the shared
scratch.  Both entry points run the same kernels with the same slicing, so the results must agree to rounding (1e-6; in
practice bit for bit).  This file sorts last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _s():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _run_both(lib, L, B, F, loss_kind, label, seed):
    g = torch.Generator().manual_seed(seed)
    x = (torch.rand((B, F), generator=g) * 2 - 1).to(torch.bfloat16).cuda()
    w = ((torch.rand((F, 1), generator=g) * 2 - 1) * 0.01).cuda()
    bias = torch.tensor([0.05], device="cuda")
    ws = torch.empty(64 * B, device="cuda")
    outs = []
    for static in (False, True):
        logits = torch.full((B, 1), float("nan"), device="cuda")
        dlogit = torch.full((B,), float("nan"), device="cuda")
        loss = torch.full((2,), float("nan"), device="cuda")
        if static:
            L.check(lib.b2g_dhead_fwd_loss_static(_p(x), _p(w), None, _p(bias), B, F, loss_kind, label, 1.0 / B,
                                                  _p(logits), _p(dlogit), _p(loss), _s()), "b2g_dhead_fwd_loss_static", 2)
        else:
            L.check(lib.b2g_dhead_fwd_loss(_p(x), _p(w), None, _p(bias), B, F, loss_kind, label, 1.0 / B, _p(ws),
                                           ws.numel(), _p(logits), _p(dlogit), _p(loss), _s()), "b2g_dhead_fwd_loss", 2)
        torch.cuda.synchronize()
        outs.append((logits.cpu(), dlogit.cpu(), loss.cpu()))
    (la, da, sa), (lb, db, sb) = outs
    assert torch.isfinite(lb).all() and torch.isfinite(db).all() and torch.isfinite(sb).all()
    assert float((la - lb).abs().max()) <= 1e-6 * (1 + float(la.abs().max()))
    assert float((da - db).abs().max()) <= 1e-6 and float((sa - sb).abs().max()) <= 1e-6
    # and against plain arithmetic (same bound as tests/test_small_kernels_gpu.py uses for the head)
    ref = x.float().cpu() @ w.cpu() + 0.05
    assert float((lb - ref).abs().max()) < 2e-3 * (1 + float(ref.abs().max()))


def test_static_scratch_head_matches_workspace_head_two_calls_in_a_row():
    from imagegeneration_b200 import lib as L
    lib = L.load()
    _run_both(lib, L, 64, 512 * 18 * 32, 0, 1.0, 11)      # the benched shape: batch 64, F = 294 912, BCE vs label 1
    _run_both(lib, L, 5, 4096, 1, 0.0, 12)                # second call: other batch, other slicing, hinge; stale scratch
    _run_both(lib, L, 64, 512 * 18 * 32, 0, 0.0, 13)


def test_static_scratch_head_rejects_oversized_batches():
    from imagegeneration_b200 import lib as L
    lib = L.load()
    B, F = 70000, 8                                        # 70 000 samples x 1 slice > 65 536 scratch floats
    x = torch.zeros((B, F), dtype=torch.bfloat16, device="cuda")
    w = torch.zeros((F, 1), device="cuda")
    bias = torch.zeros(1, device="cuda")
    logits = torch.empty((B, 1), device="cuda")
    dlogit = torch.empty((B,), device="cuda")
    loss = torch.empty(2, device="cuda")
    rc = lib.b2g_dhead_fwd_loss_static(_p(x), _p(w), None, _p(bias), B, F, 0, 1.0, 1.0 / B, _p(logits), _p(dlogit),
                                       _p(loss), _s())
    assert rc != 0 and b"static scratch holds" in lib.b2g_last_error()
