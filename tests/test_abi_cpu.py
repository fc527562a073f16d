"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/wl_b200.h declares, and fails loudly (no CPU fallback) when there is no Blackwell GPU."""
import ctypes as C
import os
import re

import pytest

import __graft_entry__ as ge

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def wl():
    ge.build()
    import waterlily_jl_b200 as wl
    return wl


def test_header_symbols_exported(wl):
    hdr = open(os.path.join(ROOT, "include", "wl_b200.h")).read()
    declared = set(re.findall(r"\b(wl_[a-z_0-9]+)\s*\(", hdr))
    declared -= {"wl_sim"}
    lib = wl.load_library()
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert set(wl.ABI_SYMBOLS) == declared, set(wl.ABI_SYMBOLS) ^ declared
    assert lib.wl_version() >= 100


def test_struct_sizes_match_header(wl):
    # POD layouts the Julia/ctypes shims mirror
    assert C.sizeof(wl.WLConfig) == 4 * (1 + 3 + 3 + 1 + 1 + 1 + 1 + 3 + 1 + 1 + 1 + 1 + 1)
    assert C.sizeof(wl.WLBody) == 4 * (2 + 3 + 2 + 9 + 1)
    assert C.sizeof(wl.WLStepInfo) == 4 * (4 + 2 + 3 * 2 * 34)


def test_no_cpu_fallback(wl):
    """Without a GPU the product path must raise, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(wl.WaterLilyB200Error) as e:
        wl.Simulation((16, 16), (1.0, 0.0))
    assert "no CPU fallback" in str(e.value) or "no CUDA device" in str(e.value)


def test_unsupported_features_are_explicit(wl):
    with pytest.raises(AssertionError, match="U must be specified"):      # src/WaterLily.jl:99
        wl.Simulation((16, 16), lambda i, x, t: 1.0)
    with pytest.raises(wl.WaterLilyB200Error, match="UNSUPPORTED"):
        wl.Simulation((16, 16), (1.0, 0.0), T="float64")
    with pytest.raises(wl.WaterLilyB200Error, match="UNSUPPORTED"):
        wl.Simulation((16, 16), (1.0, 0.0), scheme="mine")


def test_product_does_not_import_oracle():
    """The product package must not reference the oracle (test infrastructure)."""
    pkg = os.path.join(ROOT, "waterlily.jl_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".jl")):
                assert "oracle" not in open(os.path.join(dp, f), errors="ignore").read().lower(), f
