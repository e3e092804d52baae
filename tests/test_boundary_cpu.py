"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/celerite_b200.h declares (no compute calls: there is no GPU here), and the product path
fails loudly instead of falling back to the CPU."""
import ctypes
import os
import re

import pytest

from celerite_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "celerite_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cel_[a-zA-Z0-9_]+)\s*\(", text)))


def test_header_declares_something():
    syms = declared_symbols()
    assert "cel_compute" in syms and "cel_batched_log_likelihood" in syms and len(syms) >= 20


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_capi.LIB_PATH), "run python -m celerite_b200.build"
    lib = ctypes.CDLL(_capi.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_ctypes_table_matches_header():
    assert sorted(_capi.SIGNATURES) == declared_symbols()


def test_version_string():
    assert b"sm_100a" in _capi.lib().cel_version()


def test_no_cpu_fallback():
    """Without a GPU, creating a solver must raise (not silently compute on the host)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from celerite_b200 import solver
    with pytest.raises(_capi.CeleriteB200Error):
        solver.CholeskySolver()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "celerite_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "celerite_oracle" not in src and "libceleref" not in src, f
