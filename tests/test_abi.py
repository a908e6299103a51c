"""CPU: the C-ABI library loads and exports exactly what include/aha_b200.h declares (no compute calls without a GPU)."""
import ctypes
import os
import re

import pytest

from aha_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "aha_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(aha_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/aha_b200.h but not exported"


def test_binding_table_matches_header():
    assert sorted(_lib.SIGNATURES) == _declared()


def test_status_codes_and_error_string():
    lib = _lib.load()
    # argument validation happens before any CUDA call
    rc = lib.aha_find_ticks(None, 1, 10, 0.02, 10, None, None, 4, None, None, None, None)
    assert rc == -1
    with pytest.raises(ValueError, match="window_length must be less than or equal to the size of x"):
        _lib.check(rc)
    rc = lib.aha_cache_create(1, 2, 60, 0, 16, 4, 8, 0, ctypes.byref(ctypes.c_void_p()))
    assert rc == -1 and b"head_dim" in lib.aha_last_error()
    rc = lib.aha_cache_create(1, 2, 64, 0, 16, 16, 8, 0, ctypes.byref(ctypes.c_void_p()))
    assert rc == -1 and b"num_sink_tokens" in lib.aha_last_error()
    assert lib.aha_cache_destroy(None) == 0 and lib.aha_model_destroy(None) == 0


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libaha_b200.so")
    with pytest.raises(RuntimeError, match="no CPU / PyTorch fallback"):
        _lib.load()


def test_product_path_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "aha_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            assert "oracle" not in open(os.path.join(pkg, fn)).read().replace("# oracle", ""), fn
