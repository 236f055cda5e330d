"""ctypes binding of libb200gan.so (the C ABI declared in include/b200gan.h).

The product path has no CPU or library fallback: if the shared library is missing or a call fails, a
`B200GanError` is raised.  Only raw device pointers, sizes and the CUDA stream handle cross this boundary.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = PKG_DIR / "libb200gan.so"

# enums mirrored from include/b200gan.h
CONV_K3S1, CONV_K4S2, CONVT_K4S2, CONV_K1S1, CONV_K3S1_C3 = 0, 1, 2, 3, 4
ACT_NONE, ACT_LEAKY, ACT_RELU, ACT_TANH = 0, 1, 2, 3


class B200GanError(RuntimeError):
    pass


class ConvArgs(C.Structure):
    _fields_ = [
        ("kind", C.c_int),
        ("N", C.c_int), ("Hin", C.c_int), ("Win", C.c_int), ("Cin", C.c_int),
        ("Cout", C.c_int),
        ("inp", C.c_void_p),
        ("wpacked", C.c_void_p),
        ("out", C.c_void_p),
        ("bias", C.c_void_p),
        ("act", C.c_int),
        ("alpha", C.c_float),
        ("dropout_thresh8", C.c_int),
        ("dropout_seed", C.c_ulonglong),
        ("yprev", C.c_void_p),
        ("dact", C.c_int),
        ("dact_alpha", C.c_float), ("dact_scale", C.c_float), ("dact_zero", C.c_float),
        ("out_f32", C.c_int), ("cout_store", C.c_int),
        ("tf32", C.c_int),
        ("colsum", C.c_void_p),
        ("dropout_seed_dev", C.c_void_p),
        ("colsumsq", C.c_void_p),
    ]


class WgradArgs(C.Structure):
    _fields_ = [
        ("kind", C.c_int),
        ("N", C.c_int), ("Hs", C.c_int), ("Ws", C.c_int),
        ("Ca", C.c_int), ("Cb", C.c_int),
        ("shifted", C.c_void_p),
        ("unshifted", C.c_void_p),
        ("dw", C.c_void_p),
        ("ca_valid", C.c_int),
        ("dbias", C.c_void_p),
        ("tf32", C.c_int),
    ]


class PackJob(C.Structure):
    _fields_ = [
        ("kind", C.c_int), ("kh", C.c_int), ("kw", C.c_int), ("D2", C.c_int), ("D3", C.c_int),
        ("n_is_d3", C.c_int), ("flip", C.c_int), ("ck_pad", C.c_int),
        ("w", C.c_void_p), ("scale_scalar", C.c_void_p), ("scale_d2", C.c_void_p), ("scale_d3", C.c_void_p),
        ("wpacked", C.c_void_p),
        ("f32", C.c_int),
    ]


_lib = None


def load() -> C.CDLL:
    """Loads libb200gan.so once; raises if it has not been built (see imagegeneration_b200/build.py)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise B200GanError(
            f"{LIB_PATH} is missing: build it with `python -m imagegeneration_b200.build` "
            "(there is no CPU fallback for the B200 kernels)")
    lib = C.CDLL(str(LIB_PATH))
    lib.b2g_version.restype = C.c_int
    lib.b2g_last_error.restype = C.c_char_p
    lib.b2g_conv_igemm.restype = C.c_int
    lib.b2g_conv_igemm.argtypes = [C.POINTER(ConvArgs), C.c_void_p]
    lib.b2g_pack_conv_weight.restype = C.c_int
    lib.b2g_pack_conv_weight.argtypes = [C.c_int] * 8 + [C.c_void_p] * 5 + [C.c_void_p]
    lib.b2g_pack_conv_weight_multi.restype = C.c_int
    lib.b2g_pack_conv_weight_multi.argtypes = [C.POINTER(PackJob), C.c_int, C.c_void_p]
    lib.b2g_pack_conv_weight_f32.restype = C.c_int
    lib.b2g_pack_conv_weight_f32.argtypes = [C.c_int] * 8 + [C.c_void_p] * 5 + [C.c_void_p]
    lib.b2g_conv_plan_query.restype = C.c_int
    lib.b2g_conv_plan_query.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int), C.c_int]
    lib.b2g_pack_source_index.restype = C.c_long
    lib.b2g_pack_source_index.argtypes = [C.c_int] * 8 + [C.c_long]
    _declare_rest(lib)
    _lib = lib
    return lib


def _declare_rest(lib: C.CDLL) -> None:
    """argtypes for the entry points added after the conv kernels (kept tolerant while the ABI grows)."""
    for name, argtypes in _EXTRA_SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is None:
            raise B200GanError(f"libb200gan.so does not export {name}; rebuild the library")
        fn.restype = C.c_int
        fn.argtypes = argtypes


_P, _I, _F, _L, _U64 = C.c_void_p, C.c_int, C.c_float, C.c_long, C.c_ulonglong
_EXTRA_SIGNATURES: dict[str, list] = {
    "b2g_conv_wgrad": [C.POINTER(WgradArgs), _P],
    # in, out, N, H, W, stream
    "b2g_im2col27": [_P, _P, _I, _I, _I, _P],
    # in, w, wscale, bias, out, N, H, W, act, alpha, thresh8, seed, yprev, dact, stream
    "b2g_conv3x3_c3_to_c64": [_P, _P, _P, _P, _P, _I, _I, _I, _I, _F, _I, _U64, _P, _I, _P],
    # in, w, wscale, out, N, H, W, act, yprev, dact, stream
    "b2g_conv3x3_c64_to_c3": [_P, _P, _P, _P, _I, _I, _I, _I, _P, _I, _P],
    # t3, t64, dw, dbias, N, H, W, stream
    "b2g_wgrad3x3_c3_c64": [_P, _P, _P, _P, _I, _I, _I, _P],
    # z, W, gamma, beta, inv_std, out, B, F, stream
    "b2g_dense_g_fwd": [_P, _P, _P, _P, _F, _P, _I, _I, _P],
    # z, W, gamma, inv_std, dq, dW, dgamma, dbeta, B, F, stream
    "b2g_dense_g_bwd": [_P, _P, _P, _F, _P, _P, _P, _P, _I, _I, _P],
    # x, w, wscale, bias, B, F, loss_kind, label, inv_batch, ws, ws_floats, logits, dlogit, loss_out, stream
    "b2g_dhead_fwd_loss": [_P, _P, _P, _P, _I, _I, _I, _F, _F, _P, _I, _P, _P, _P, _P],
    # x, w, wscale, bias, B, F, loss_kind, label, inv_batch, logits, dlogit, loss_out, stream  (static scratch, no alloc)
    "b2g_dhead_fwd_loss_static": [_P, _P, _P, _P, _I, _I, _I, _F, _F, _P, _P, _P, _P],
    # x, w, wscale, dlogit, B, F, alpha, dscale, dzero, dx, dw, db, stream
    "b2g_dhead_bwd": [_P, _P, _P, _P, _I, _I, _F, _F, _F, _P, _P, _P, _P],
    # w, g, m, v, n, lr_t, b1, b2, eps, gscale, stream
    "b2g_adam": [_P, _P, _P, _P, _L, _F, _F, _F, _F, _F, _P],
    # w, g, m, v, n, lr_t_dev, b1, b2, eps, gscale, stream
    "b2g_adam_dev": [_P, _P, _P, _P, _L, _P, _F, _F, _F, _F, _P],
    "b2g_adam_nvls": [_P, _P, _P, _P, _P, _L, _L, _F, _P, _F, _F, _F, _F, _P],
    # step_ctr, seed_base, seeds21, lr3, lr_g, lr_d, b1, b2, stream
    "b2g_step_prologue": [_P, C.c_ulonglong, _P, _P, _F, _F, _F, _F, _P],
    # x, out, P, C, stream
    "b2g_colsum_bf16": [_P, _P, _L, _I, _P],
    # W, G, gamma, inv_std, dgamma, taps, Co, Ci, stream
    "b2g_bn_fold_finalize": [_P, _P, _P, _F, _P, _I, _I, _I, _P],
    # n, W**, u**, v_raw**, t_raw**, scal**, K*, C*, stream
    "b2g_spectral_norm_multi": [_I, _P, _P, _P, _P, _P, _P, _P, _P],
    # W, G, u, v_raw, scal, dot_scratch, K, C, stream
    "b2g_sn_grad_finalize": [_P, _P, _P, _P, _P, _P, _I, _I, _P],
    # w, g, n_layers, off*, K*, C*, dots, stream
    "b2g_sn_dot_multi": [_P, _P, _I, _P, _P, _P, _P, _P],
    # w, g, m, v, n, n_layers, off*, K*, C*, u**, v_raw**, t_raw**, scal**, dots, lr_t, lr_t_dev, b1, b2, eps, gscale, stream
    "b2g_adam_sn": [_P, _P, _P, _P, _L, _I, _P, _P, _P, _P, _P, _P, _P, _P, _F, _P, _F, _F, _F, _F, _P],
    # fp32 variants of the edge / head / stem kernels (the fp32/tf32 parity path)
    "b2g_conv3x3_c3_to_c64_f32": [_P, _P, _P, _P, _P, _I, _I, _I, _I, _F, _I, _U64, _P, _I, _P],
    "b2g_conv3x3_c64_to_c3_f32": [_P, _P, _P, _P, _I, _I, _I, _I, _P, _I, _P],
    "b2g_wgrad3x3_c3_c64_f32": [_P, _P, _P, _P, _I, _I, _I, _P],
    "b2g_dhead_fwd_loss_f32": [_P, _P, _P, _P, _I, _I, _I, _F, _F, _P, _I, _P, _P, _P, _P],
    "b2g_dhead_bwd_f32": [_P, _P, _P, _P, _I, _I, _F, _F, _F, _P, _P, _P, _P],
    "b2g_dense_g_fwd_f32": [_P, _P, _P, _P, _F, _P, _I, _I, _P],
    # z, W, gamma, inv_std, dq, scratch, dW, dgamma, dbeta, B, F, stream
    "b2g_dense_g_bwd_f32": [_P, _P, _P, _F, _P, _P, _P, _P, _P, _I, _I, _P],
    # CycleGAN path (csrc/b2g_cyclegan.cu)
    # in, in_f32, N, Hin, Win, C, ldc, kh, kw, mode, stride, pad_t, pad_l, reflect, Hout, Wout, out, Mpad, Kpad, stream
    "b2g_im2col": [_P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P, _L, _I, _P],
    # w, taps, D2, D3, n_is_d3, out, Npad, Kpad, stream
    "b2g_pack_gemm_weight": [_P, _I, _I, _I, _I, _P, _I, _I, _P],
    # dpad, dx, N, H, W, C, ldc, pad, stream
    "b2g_reflect_fold": [_P, _P, _I, _I, _I, _I, _I, _I, _P],
    # a, res, gamma, beta, y, stat, N, H, W, C, ldc, act, alpha, eps, y_f32, stream
    "b2g_rownorm_fwd": [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _F, _I, _P],
    # dy, y, a, res, gamma, stat, da, dgamma, dbeta, N, H, W, C, ldc, act, alpha, y_f32, stream
    "b2g_rownorm_bwd": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _I, _P],
    # a, b, n, scale, loss_out, da, stream
    "b2g_l1_loss": [_P, _P, _L, _F, _P, _P, _P],
    # g, y, out, n, f32, kind, alpha, scale, zero, stream
    "b2g_dact": [_P, _P, _P, _L, _I, _I, _F, _F, _F, _P],
    # logits, labels, B, loss_out, dlogit, stream
    "b2g_bce_logits": [_P, _P, _I, _P, _P, _P],
    # images_u8, idx, out, B, per_image, stream
    "b2g_gather_u8_to_f32": [_P, _P, _P, _I, _L, _P],
    # x, out, P, C, stream
    "b2g_colsum_f32": [_P, _P, _L, _I, _P],
    # z, W, gamma, beta, inv_std, alpha, out, B, F, stream
    "b2g_dense_g_fwd_leaky": [_P, _P, _P, _P, _F, _F, _P, _I, _I, _P],
    # x, sums, P, C, stream
    "b2g_bn_stats": [_P, _P, _L, _I, _P],
    # sums, P, C, eps, momentum, stat, moving_mean, moving_var, bias, stream
    "b2g_bn_finalize": [_P, _L, _I, _F, _F, _P, _P, _P, _P, _P],
    # moving_mean, moving_var, C, eps, stat, bias, stream
    "b2g_bn_inference_stat": [_P, _P, _I, _F, _P, _P, _P],
    # x, stat, gamma, beta, alpha, y, P, C, stream
    "b2g_bn_apply_act": [_P, _P, _P, _P, _F, _P, _L, _I, _P],
    # g, x, stat, gamma, red, dx, dgamma, dbeta, P, C, stream
    "b2g_bn_bwd": [_P, _P, _P, _P, _P, _P, _P, _P, _L, _I, _P],
    # w, g, rms, n, lr, rho, eps, gscale, clip_n, clip, stream
    "b2g_rmsprop": [_P, _P, _P, _L, _F, _F, _F, _F, _L, _F, _P],
    # n
    "b2g_set_sm_limit": [_I],
    "b2g_num_sms": [],
    # x, out, N, H, W, C, stream
    "b2g_avgpool8_bf16": [_P, _P, _I, _I, _I, _I, _P],
    # kind, N, Hin, Win, Cin, block_n, phase_group, out, out_len   (host-only)
    "b2g_conv_rows_plan_query": [_I, _I, _I, _I, _I, _I, _I, C.POINTER(C.c_int), _I],
    "b2g_conv_rows_plan_query_merged": [_I, _I, _I, _I, _I, _I, _I, _I, C.POINTER(C.c_int), _I],
}


profile_sink = None  # when a list: ops.* append (kernel, flops, start_event, end_event) around tensor-core launches
launch_count = 0  # kernels launched through this binding (bench.py reports it as gpu_launches)


def check(rc: int, what: str, launches: int = 1) -> None:
    global launch_count
    launch_count += launches
    if rc != 0:
        msg = load().b2g_last_error()
        raise B200GanError(f"{what} failed (code {rc}): {msg.decode() if msg else '?'}")


# Names every build of the library must export (tests/test_abi.py checks them against include/b200gan.h).
def exported_symbols() -> list[str]:
    return ["b2g_version", "b2g_last_error", "b2g_conv_igemm", "b2g_pack_conv_weight", "b2g_pack_conv_weight_f32",
            "b2g_pack_conv_weight_multi",
            "b2g_conv_plan_query",
            "b2g_pack_source_index", *_EXTRA_SIGNATURES]
