"""CPU: the C-ABI library loads and exports every symbol include/b200k.h declares (no compute calls)."""

import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200k.h")


def _declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200k_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_entry_points():
    syms = _declared_symbols()
    assert "b200k_rmsnorm_fwd" in syms and "b200k_version" in syms


def test_library_exports_every_declared_symbol():
    from b200ex import _lib

    path = _lib.library_path()
    assert os.path.exists(path), f"{path} missing: run `make -C lightning-thunder_b200/csrc`"
    handle = ctypes.CDLL(path)
    for s in _declared_symbols():
        assert hasattr(handle, s), f"{s} declared in include/b200k.h but not exported by libb200k.so"


def test_binding_table_matches_header():
    from b200ex import _lib

    assert sorted(_lib.EXPORTS) == _declared_symbols()
    L = _lib.lib()
    assert L.b200k_version() == _lib.ABI_VERSION
    assert L.b200k_error_string(0) == b"success"
    assert b"scratch" in L.b200k_error_string(-4)


def test_ops_refuse_cpu_tensors():
    import torch

    from b200ex import ops

    x = torch.zeros(2, 8)
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        ops.rmsnorm_fwd(x, None, None, 1e-5)


def test_missing_library_fails_loudly(monkeypatch):
    from b200ex import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setenv("B200K_LIBRARY", "/nonexistent/libb200k.so")
    with pytest.raises(_lib.B200LibraryError):
        _lib.lib()
