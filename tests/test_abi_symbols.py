"""CPU: the C-ABI library loads and exports every symbol include/ssvi_b200.h declares (no compute)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "ssvi_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ssvi_[a-z_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    from ssvi_tf_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), "missing export: " + n
        assert n in _lib.SIGNATURES, "ctypes binding missing for " + n
    assert set(_lib.SIGNATURES) == set(names)


def test_every_entry_point_is_documented_for_the_reference_maintainer():
    """INTEGRATION.md's table (entry point <-> the reference lines it replaces) names every symbol of the header."""
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    missing = [n for n in _declared() if n not in doc]
    assert not missing, missing


def test_version_and_acc_len():
    from ssvi_tf_b200 import _lib
    lib = _lib.load()
    assert b"sm_100a" in lib.ssvi_version()
    c = _lib.SsviConfig()
    c.rank, c.cov_kind = 20, _lib.COV_FULL
    assert lib.ssvi_acc_len(ctypes.byref(c)) == 20 + 400 + 2
    c.cov_kind = _lib.COV_DIAG
    assert lib.ssvi_acc_len(ctypes.byref(c)) == 20 + 20 + 2


def test_config_struct_matches_header_layout():
    from ssvi_tf_b200 import _lib
    # 12 int32 + 7 double + 8 int64
    assert ctypes.sizeof(_lib.SsviConfig) == 14 * 4 + 7 * 8 + 8 * 8


def test_create_rejects_bad_config_without_gpu():
    from ssvi_tf_b200 import _lib
    lib = _lib.load()
    c = _lib.SsviConfig()
    c.order, c.rank, c.k1, c.k2 = 1, 4, 4, 4          # order < 2
    h = ctypes.c_void_p()
    assert lib.ssvi_create(ctypes.byref(c), ctypes.byref(h)) == -1
    assert b"order" in lib.ssvi_last_error(None)
    c.order, c.mean_update = 3, 7
    for d in range(3):
        c.dims[d] = 4
    assert lib.ssvi_create(ctypes.byref(c), ctypes.byref(h)) == -1
    assert b"Unidentified update formula" in lib.ssvi_last_error(None)


def test_engine_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from ssvi_tf_b200.engine import Engine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Engine(dict(model="d", likelihood="normal", D=4, diag=False, k1=4, k2=4, batch_size=4, mean_update="S",
                    cov_update="N", eta=1, cov_eta=1, sigma_eta=1), [3, 3, 3], [[0, 0, 0]], [1.0])


def test_enum_constants_match_the_header():
    """The ctypes mirror (_lib.py) carries the header's enumerators by value: check the ones the host code switches on."""
    from ssvi_tf_b200 import _lib
    src = open(os.path.join(ROOT, "include", "ssvi_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    vals = {k: int(v) for k, v in re.findall(r"\b(SSVI_[A-Z0-9_]+)\s*=\s*(-?\d+)", src)}
    want = {"SSVI_TABLE_F64": _lib.TABLE_F64, "SSVI_TABLE_F32": _lib.TABLE_F32, "SSVI_TABLE_F32_TC": _lib.TABLE_F32_TC,
            "SSVI_FLAG_NONFINITE_ACC": _lib.FLAG_NONFINITE_ACC, "SSVI_FLAG_NONFINITE_COV": _lib.FLAG_NONFINITE_COV,
            "SSVI_FLAG_PHI_ZERO": _lib.FLAG_PHI_ZERO, "SSVI_FLAG_CHOL_FAIL": _lib.FLAG_CHOL_FAIL,
            "SSVI_FLAG_EIG_NOCONV": _lib.FLAG_EIG_NOCONV, "SSVI_FLAG_TC_RANGE": _lib.FLAG_TC_RANGE}
    for name, v in want.items():
        assert vals[name] == v, name
    tune = [vals[k] for k in vals if k.startswith("SSVI_TUNE_")]
    assert len(tune) == len(set(tune)) and all(t & (t - 1) == 0 for t in tune)      # distinct single bits


def test_create_rejects_bad_fp32_option_fields():
    from ssvi_tf_b200 import _lib
    lib = _lib.load()
    c = _lib.SsviConfig()
    c.order, c.rank, c.k1, c.k2 = 3, 8, 64, 4
    c.mean_update = c.cov_update = _lib.UPDATE_S
    for d in range(3):
        c.dims[d] = 4
    h = ctypes.c_void_p()
    c.table_dtype = 3
    assert lib.ssvi_create(ctypes.byref(c), ctypes.byref(h)) == -1
    assert b"table_dtype" in lib.ssvi_last_error(None)
    c.table_dtype, c.tc_scale_log2 = _lib.TABLE_F32_TC, 15
    assert lib.ssvi_create(ctypes.byref(c), ctypes.byref(h)) == -1
    assert b"tc_scale_log2" in lib.ssvi_last_error(None)
