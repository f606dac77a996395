"""The C-ABI library loads on a CPU-only box, exports every symbol include/cgb200.h declares, and
refuses to compute without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

from conjugategradient_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "cgb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cgb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported():
    lib = L.load()
    names = declared_symbols()
    assert len(names) >= 50
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/cgb200.h but not exported"


def test_binding_table_matches_header():
    assert sorted(L.SIGNATURES) == declared_symbols()


def test_struct_layouts_match_header():
    # cgb_cg_opts: 12 ints; cgb_cg_stats layout mirrors the header field order
    assert C.sizeof(L.CgOpts) == 48
    assert C.sizeof(L.CgStats) == 8 + 8 + 8 + 4 + 4 + 8 * 4 + 8 + 8 + 4 + 4 + 32


def test_footprint_formula():
    # reference src/CG.hpp:555-558
    assert L.load().cgb_memory_footprint(16384, 81408) == 1957888


@pytest.mark.skipif(L.load().cgb_device_count() > 0, reason="GPU present")
def test_no_cpu_fallback():
    lib = L.load()
    h = C.c_void_p()
    rc = lib.cgb_ctx_create(0, C.byref(h))
    assert rc == L.CGB_ERR_CUDA and not h.value
    assert b"no CPU fallback" in lib.cgb_last_error()
    from conjugategradient_b200 import solver
    with pytest.raises(L.CgbError):
        solver.Context(0)
