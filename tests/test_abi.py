"""The C-ABI library loads and exports every symbol include/qdc_b200.h declares, and the ctypes
signature table covers exactly that set (no compute calls here: no GPU in this container)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "qdc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(qdc_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported_and_bound():
    from qdc_b200 import _lib
    names = _declared()
    assert "qdc_contract_f64" in names and "qdc_mis_reduce" in names
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in names:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert sorted(_lib.SIGNATURES) == names
    assert _lib.load().qdc_version() >= 100


def test_no_cpu_fallback(monkeypatch):
    """A missing extension must raise, never silently compute on the CPU."""
    from qdc_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libqdc_b200.so")
    with pytest.raises(_lib.QdcError):
        _lib.load()


def test_product_does_not_import_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py may touch oracle/."""
    pkg = os.path.join(ROOT, "dqva-and-circuit-cutting_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), fn
                assert "liboracle" not in src, fn
