"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports
every symbol include/agents_b200.h declares, and refuses to run without a GPU
(no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

import agents_jl_b200 as ab
from agents_jl_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "agents_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ab_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert _header_symbols() == ab.declared_symbols()


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(L.LIB_PATH)
    for name in _header_symbols():
        assert hasattr(lib, name), f"{name} not exported"
    assert ab.lib().ab_version() >= 100


def test_struct_layouts_match_header():
    # sizes implied by include/agents_b200.h (natural alignment)
    assert C.sizeof(L.SpaceDesc) == 4 + 4 + 24 + 8 + 3 + 5 + 4 + 4 + 24
    assert C.sizeof(L.FlockParams) == 6 * 8 + 4 * 4
    assert C.sizeof(L.PairForceParams) == 2 * 8 + 4 * 4
    assert C.sizeof(L.SchellingParams) == 16


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(ab.AgentsError) as ei:
        ab.StandardABM(ab.ContinuousSpace((10.0, 10.0), spacing=1.0))
    assert ei.value.code == L.ERR_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "agents.jl_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.lower() or f == "__init__.py", f"{f} mentions the oracle"
