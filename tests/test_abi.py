"""CPU: libb200gan.so builds/loads and exports every symbol include/b200gan.h declares (no compute calls)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200gan.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2g_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_families():
    names = _declared_symbols()
    for fam in ("b2g_conv_igemm", "b2g_conv_wgrad", "b2g_pack_conv_weight", "b2g_dense_g_fwd", "b2g_dense_g_bwd",
                "b2g_dhead_fwd_loss", "b2g_dhead_bwd", "b2g_adam", "b2g_spectral_norm_multi", "b2g_im2col27",
                "b2g_last_error", "b2g_version"):
        assert fam in names


def test_library_exports_every_declared_symbol():
    from imagegeneration_b200 import build, lib
    path = build.build_library()
    dll = ctypes.CDLL(str(path))
    for name in _declared_symbols():
        assert hasattr(dll, name), f"{name} declared in include/b200gan.h but not exported"
    loaded = lib.load()
    for name in lib.exported_symbols():
        assert hasattr(loaded, name)
    assert loaded.b2g_version() >= 100


def test_binding_covers_the_header():
    from imagegeneration_b200 import lib
    assert sorted(lib.exported_symbols()) == _declared_symbols()


def test_errors_are_reported_not_thrown():
    """Invalid arguments come back as negative codes + message (never a crash), also without a GPU."""
    from imagegeneration_b200 import lib
    l = lib.load()
    out = (ctypes.c_int * 80)()
    assert l.b2g_conv_plan_query(0, 1, 18, 48, 128, out, 80) == -2      # width 48 is not tileable
    assert b"cannot tile" in l.b2g_last_error()
    assert l.b2g_conv_plan_query(0, 1, 18, 32, 128, out, 10) == -1      # output buffer too small
    assert l.b2g_conv_igemm(None, None) == -1


def test_product_path_has_no_cpu_fallback():
    import pytest
    import torch
    from imagegeneration_b200 import ops
    from imagegeneration_b200.lib import B200GanError
    with pytest.raises(B200GanError):
        ops.im2col27(torch.zeros(1, 8, 8, 3))          # CPU tensor -> loud failure, not a silent fallback
    src = open(os.path.join(ROOT, "imagegeneration_b200", "sndcgan.py")).read() + \
        open(os.path.join(ROOT, "imagegeneration_b200", "ops.py")).read()
    assert "oracle" not in src.replace("oracle/dropout_hash.py", "")   # the product never imports the oracle
