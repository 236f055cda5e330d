"""Tensor-level wrappers over the C ABI: they validate shapes, allocate outputs with torch (device memory and
streams are torch's job) and pass raw pointers to libb200gan.so.  No arithmetic happens in Python/torch here.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lib as _lib
from .lib import WgradArgs
from .lib import (ACT_LEAKY, ACT_NONE, ACT_RELU, ACT_TANH, CONV_K1S1, CONV_K3S1, CONV_K3S1_C3, CONV_K4S2, CONVT_K4S2,
                  ConvArgs,
                  B200GanError, check)

__all__ = ["pack_conv_weight", "conv_igemm", "conv_wgrad", "im2col27", "CONV_K1S1", "CONV_K3S1_C3", "CONV_K3S1", "CONV_K4S2", "CONVT_K4S2", "ACT_NONE", "ACT_LEAKY",
           "ACT_RELU", "ACT_TANH"]

_KIND_GEOM = {CONV_K3S1: (3, 1, 9), CONV_K4S2: (4, 1, 16), CONVT_K4S2: (4, 4, 4), CONV_K1S1: (1, 1, 1),
              CONV_K3S1_C3: (1, 1, 1)}  # k, phases, taps (K3S1_C3 takes the K1S1 packing of the [27][Cout] matrix)


def _ptr(t: torch.Tensor | None):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _prof_begin():
    if _lib.profile_sink is None:
        return None
    ev = torch.cuda.Event(enable_timing=True)
    ev.record()
    return ev


def _prof_end(ev, name, flops, label=""):
    if ev is None:
        return
    end = torch.cuda.Event(enable_timing=True)
    end.record()
    _lib.profile_sink.append((name, flops, ev, end, label))


def _need_cuda(*ts: torch.Tensor | None) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise B200GanError("b200gan ops need CUDA tensors: there is no CPU fallback")
        if t is not None and not t.is_contiguous():
            raise B200GanError("b200gan ops need contiguous tensors")


def pack_conv_weight(kind: int, w: torch.Tensor, n_is_d3: bool, flip: bool = False,
                     scale_scalar: torch.Tensor | None = None, scale_d2: torch.Tensor | None = None,
                     scale_d3: torch.Tensor | None = None, out: torch.Tensor | None = None,
                     ck_pad: int = 0, tf32: bool = False) -> torch.Tensor:
    """fp32 Keras-layout conv weight [kh,kw,d2,d3] -> bf16 packed [phases, Cn, taps*Ck] for conv_igemm
    (tf32=True: fp32 packing for the kind::tf32 path)."""
    _need_cuda(w, scale_scalar, scale_d2, scale_d3, out)
    k, phases, taps = _KIND_GEOM[kind]
    kh, kw, d2, d3 = w.shape
    if w.dtype != torch.float32 or kh != k or kw != k:
        raise B200GanError(f"pack_conv_weight: expected fp32 [{k},{k},d2,d3], got {tuple(w.shape)} {w.dtype}")
    cn, ck = (d3, d2) if n_is_d3 else (d2, d3)
    ck = max(ck, ck_pad)
    odt = torch.float32 if tf32 else torch.bfloat16
    if out is None:
        out = torch.empty((phases, cn, taps * ck), dtype=odt, device=w.device)
    elif out.dtype != odt:
        raise B200GanError("pack_conv_weight: output dtype does not match the path")
    fn = _lib.load().b2g_pack_conv_weight_f32 if tf32 else _lib.load().b2g_pack_conv_weight
    check(fn(kind, kh, kw, d2, d3, int(n_is_d3), int(flip), ck_pad, _ptr(w), _ptr(scale_scalar), _ptr(scale_d2),
             _ptr(scale_d3), _ptr(out), _stream()), "b2g_pack_conv_weight")
    return out


class PackPlan:
    """All bf16 weight packings of one model, run as ONE launch (b2g_pack_conv_weight_multi).

    add() takes the arguments of pack_conv_weight, allocates the packed tensor and returns it; run() re-packs every
    job from the current master weights.  The tensors handed to add() must stay alive and in place (arena views)."""

    MAX_JOBS = 24

    def __init__(self):
        self._jobs = []
        self._keep = []
        self._arr = None

    def add(self, kind: int, w: torch.Tensor, n_is_d3: bool, flip: bool = False,
            scale_scalar: torch.Tensor | None = None, scale_d2: torch.Tensor | None = None,
            scale_d3: torch.Tensor | None = None, ck_pad: int = 0, tf32: bool = False) -> torch.Tensor:
        """tf32=True: fp32 packing rounded to tf32 (the kind::tf32 conv path) instead of bf16."""
        _need_cuda(w, scale_scalar, scale_d2, scale_d3)
        k, phases, taps = _KIND_GEOM[kind]
        kh, kw, d2, d3 = w.shape
        if w.dtype != torch.float32 or kh != k or kw != k or not w.is_contiguous():
            raise B200GanError(f"PackPlan.add: expected contiguous fp32 [{k},{k},d2,d3], got {tuple(w.shape)} {w.dtype}")
        if len(self._jobs) >= self.MAX_JOBS:
            raise B200GanError(f"PackPlan: more than {self.MAX_JOBS} jobs")
        cn, ck = (d3, d2) if n_is_d3 else (d2, d3)
        out = torch.empty((phases, cn, taps * max(ck, ck_pad)), dtype=torch.float32 if tf32 else torch.bfloat16,
                          device=w.device)
        self._jobs.append(_lib.PackJob(kind=kind, kh=kh, kw=kw, D2=d2, D3=d3, n_is_d3=int(n_is_d3), flip=int(flip),
                                       ck_pad=ck_pad, w=_ptr(w), scale_scalar=_ptr(scale_scalar),
                                       scale_d2=_ptr(scale_d2), scale_d3=_ptr(scale_d3), wpacked=_ptr(out),
                                       f32=int(tf32)))
        self._keep.append((w, scale_scalar, scale_d2, scale_d3, out))
        self._arr = None
        return out

    def run(self) -> None:
        if not self._jobs:
            return
        if self._arr is None:
            self._arr = (_lib.PackJob * len(self._jobs))(*self._jobs)
        check(_lib.load().b2g_pack_conv_weight_multi(self._arr, len(self._jobs), _stream()),
              "b2g_pack_conv_weight_multi")


def conv_igemm(kind: int, x: torch.Tensor, wpacked: torch.Tensor, out: torch.Tensor | None = None, *,
               bias: torch.Tensor | None = None, act: int = ACT_NONE, alpha: float = 0.0,
               dropout_thresh8: int = 0, dropout_seed: int = 0, yprev: torch.Tensor | None = None,
               dact: int = ACT_NONE, dact_alpha: float = 0.0, dact_scale: float = 1.0,
               dact_zero: float = 0.0, out_f32: bool = False, tf32: bool = False,
               colsum: torch.Tensor | None = None, colsumsq: torch.Tensor | None = None) -> torch.Tensor:
    """Tensor-core implicit-GEMM conv / transposed conv with fused epilogue (see include/b200gan.h).

    out_f32=True: narrow fp32 output — wpacked has `cout_store` (<= 8) rows and out is fp32 [N,Ho,Wo,cout_store].
    colsum: fp32 [Cout], += per-channel sums of the stored outputs (bias / beta gradient fused into a backward call).
    colsumsq: fp32 [Cout], with colsum on a call without bias / dropout (forward allowed): += sums of squares — the batch
    statistics of a training-mode BatchNorm behind the conv (WGAN path).
    dropout_seed: an int, or a 1-element int64 CUDA tensor the kernel reads the seed from (CUDA-graph replay)."""
    _need_cuda(x, wpacked, out, bias, yprev, colsum, colsumsq)
    n, h, w, cin = x.shape
    k, phases, taps = _KIND_GEOM[kind]
    if kind == CONV_K3S1_C3:
        # 3-channel fp32 input, im2col patch built inside the kernel; weights = K1S1 packing of [27][Cout], ck_pad 64
        if x.dtype != torch.float32 or cin != 3 or wpacked.dtype != torch.bfloat16 or tuple(wpacked.shape[::2]) != (1, 64):
            raise B200GanError("conv_igemm(K3S1_C3): expected fp32 [N,H,W,3] and packed bf16 weights [1,Cout,64]")
    elif tf32:
        if x.dtype != torch.float32 or wpacked.dtype != torch.float32 or out_f32 or cin % 32:
            raise B200GanError("conv_igemm(tf32): fp32 activations / fp32 packed weights, Cin % 32 == 0")
        if wpacked.shape[0] != phases or wpacked.shape[2] != taps * cin:
            raise B200GanError(f"conv_igemm: packed weight {tuple(wpacked.shape)} does not match kind/Cin={cin}")
    elif x.dtype != torch.bfloat16 or wpacked.dtype != torch.bfloat16:
        raise B200GanError("conv_igemm: activations and packed weights must be bf16")
    elif wpacked.shape[0] != phases or wpacked.shape[2] != taps * cin:
        raise B200GanError(f"conv_igemm: packed weight {tuple(wpacked.shape)} does not match kind/Cin={cin}")
    cout = wpacked.shape[1]
    ho, wo = {CONV_K3S1: (h, w), CONV_K4S2: (h // 2, w // 2), CONVT_K4S2: (h * 2, w * 2), CONV_K1S1: (h, w),
              CONV_K3S1_C3: (h, w)}[kind]
    odt = torch.float32 if (out_f32 or tf32) else torch.bfloat16
    if out is None:
        out = torch.empty((n, ho, wo, cout), dtype=odt, device=x.device)
    elif tuple(out.shape) != (n, ho, wo, cout) or out.dtype != odt:
        raise B200GanError("conv_igemm: bad output tensor")
    if yprev is not None and (tuple(yprev.shape) != tuple(out.shape) or yprev.dtype != odt):
        raise B200GanError("conv_igemm: yprev must match the output")
    if bias is not None and (bias.dtype != torch.float32 or bias.numel() != cout):
        raise B200GanError("conv_igemm: bias must be fp32 [Cout]")
    if colsum is not None and (colsum.dtype != torch.float32 or colsum.numel() != cout or out_f32 or tf32
                               or bias is not None or (dact == ACT_NONE and colsumsq is None)):
        raise B200GanError("conv_igemm: colsum needs fp32 [Cout], bf16 output, no bias and a backward (dact) epilogue")
    if colsumsq is not None and (colsum is None or colsumsq.dtype != torch.float32 or colsumsq.numel() != cout
                                 or dropout_thresh8):
        raise B200GanError("conv_igemm: colsumsq needs colsum, fp32 [Cout] and no dropout")
    cout_store = 0
    if out_f32:
        cout_store, cout = cout, 32
    seed_dev = None
    if isinstance(dropout_seed, torch.Tensor):
        if dropout_seed.dtype != torch.int64 or dropout_seed.numel() != 1 or not dropout_seed.is_cuda:
            raise B200GanError("conv_igemm: a device dropout seed must be a 1-element int64 CUDA tensor")
        seed_dev, dropout_seed = dropout_seed.data_ptr(), 0
    a = ConvArgs(kind=kind, N=n, Hin=h, Win=w, Cin=cin, Cout=cout, inp=x.data_ptr(), wpacked=wpacked.data_ptr(),
                 out=out.data_ptr(), bias=None if bias is None else bias.data_ptr(), act=act, alpha=alpha,
                 dropout_thresh8=dropout_thresh8, dropout_seed=dropout_seed & 0xFFFFFFFFFFFFFFFF,
                 yprev=None if yprev is None else yprev.data_ptr(), dact=dact, dact_alpha=dact_alpha,
                 dact_scale=dact_scale, dact_zero=dact_zero, out_f32=int(out_f32), cout_store=cout_store,
                 tf32=int(tf32), colsum=None if colsum is None else colsum.data_ptr(), dropout_seed_dev=seed_dev,
                 colsumsq=None if colsumsq is None else colsumsq.data_ptr())
    ev = _prof_begin()
    check(_lib.load().b2g_conv_igemm(C.byref(a), _stream()), "b2g_conv_igemm")
    cin_alg = 27 if kind in (CONV_K1S1, CONV_K3S1_C3) else cin   # edge layers: 27 real taps*channels
    _prof_end(ev, "conv_igemm", 2.0 * n * h * w * cin_alg * (cout_store or cout) *
              (16 if kind == CONVT_K4S2 else taps) / (4 if kind == CONV_K4S2 else 1),
              f"kind{kind} {n}x{h}x{w}x{cin}->{cout_store or cout}")
    return out


def conv_wgrad(kind: int, shifted: torch.Tensor, unshifted: torch.Tensor, dw: torch.Tensor, ca_valid: int = 0,
               dbias: torch.Tensor | None = None, tf32: bool = False) -> torch.Tensor:
    """dw[tap][a][b] += sum_pixels shifted[tap-shifted pixel, a] * unshifted[pixel, b] (see include/b200gan.h).

    dw is fp32 [k,k,Ca,Cb] and is accumulated into (zero it first for a fresh gradient).  tf32=True takes fp32
    operands (kind::tf32 products, the per-layer 1e-3 parity path); the training step uses bf16 operands."""
    _need_cuda(shifted, unshifted, dw, dbias)
    n, hs, ws, ca = shifted.shape
    cb = unshifted.shape[3]
    if kind == CONV_K3S1_C3:
        # 3-channel edge layers: fp32 [N,H,W,3] x bf16 [N,H,W,64] -> dw [27][64] (+ dbias [64]); the im2col operand
        # of the K1S1 route is built inside the kernel
        if (shifted.dtype != torch.float32 or ca != 3 or unshifted.dtype != torch.bfloat16 or cb != 64
                or tuple(unshifted.shape[:3]) != (n, hs, ws) or dw.dtype != torch.float32 or dw.numel() != 27 * 64
                or tf32 or ca_valid not in (0, 27)):
            raise B200GanError("conv_wgrad(K3S1_C3): expected fp32 [N,H,W,3], bf16 [N,H,W,64] and fp32 dw [27][64]")
        a = WgradArgs(kind=kind, N=n, Hs=hs, Ws=ws, Ca=3, Cb=64, shifted=shifted.data_ptr(),
                      unshifted=unshifted.data_ptr(), dw=dw.data_ptr(), ca_valid=27,
                      dbias=None if dbias is None else dbias.data_ptr(), tf32=0)
        ev = _prof_begin()
        check(_lib.load().b2g_conv_wgrad(C.byref(a), _stream()), "b2g_conv_wgrad")
        _prof_end(ev, "conv_wgrad", 2.0 * n * hs * ws * 27 * 64, f"kind{kind} {n}x{hs}x{ws} Ca3 Cb64")
        return dw
    k = {CONV_K3S1: 3, CONV_K4S2: 4, CONV_K1S1: 1}[kind]
    hg, wg = (hs // 2, ws // 2) if kind == CONV_K4S2 else (hs, ws)
    if tuple(unshifted.shape) != (n, hg, wg, cb):
        raise B200GanError(f"conv_wgrad: unshifted operand {tuple(unshifted.shape)} != {(n, hg, wg, cb)}")
    want = torch.float32 if tf32 else torch.bfloat16
    if shifted.dtype != want or unshifted.dtype != want or dw.dtype != torch.float32:
        raise B200GanError(f"conv_wgrad: operands must be {want} and dw fp32")
    if dw.numel() != k * k * (ca_valid or ca) * cb:
        raise B200GanError("conv_wgrad: bad dw size")
    a = WgradArgs(kind=kind, N=n, Hs=hs, Ws=ws, Ca=ca, Cb=cb, shifted=shifted.data_ptr(),
                  unshifted=unshifted.data_ptr(), dw=dw.data_ptr(), ca_valid=ca_valid,
                  dbias=None if dbias is None else dbias.data_ptr(), tf32=1 if tf32 else 0)
    ev = _prof_begin()
    check(_lib.load().b2g_conv_wgrad(C.byref(a), _stream()), "b2g_conv_wgrad")
    _prof_end(ev, "conv_wgrad", 2.0 * n * hg * wg * (ca_valid or ca) * cb * k * k,
              f"kind{kind} {n}x{hs}x{ws} Ca{ca} Cb{cb}")
    return dw


def im2col27(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """fp32 [N,H,W,3] -> bf16 [N,H,W,64]: the 27 taps*channels of a 3x3 window, a ones channel, zero padding."""
    _need_cuda(x, out)
    n, h, w, c = x.shape
    if c != 3 or x.dtype != torch.float32:
        raise B200GanError("im2col27: expected fp32 [N,H,W,3]")
    if out is None:
        out = torch.empty((n, h, w, 64), dtype=torch.bfloat16, device=x.device)
    check(_lib.load().b2g_im2col27(_ptr(x), _ptr(out), n, h, w, _stream()), "b2g_im2col27")
    return out
