"""CPU: the C-ABI library loads, exports every symbol include/ro_cuda.h declares,
and its host-side tables match the oracle.  No compute calls (no GPU here)."""
import os
import re

import numpy as np
import pytest

import radio_observer_b200 as ro
from oracle import ref_numpy as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "ro_cuda.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ro_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported(lib):
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), s
    assert set(lib._ro_protos) == set(syms)


def test_struct_layouts():
    import ctypes as C
    assert C.sizeof(ro.DetRec) == 16
    assert C.sizeof(ro.WfConfig) == 4 * 4 + 8 + 4 * 2 + 4 * 8 + 4 * 5 + 4 * 2 + 4 * 6 + 4  # incl. alignment pad
    assert ro.capi.DET_DTYPE.itemsize == 16


@pytest.mark.parametrize("bins", [1024, 2048, 16384, 32768, 65536])
def test_window_table_is_the_reference_table(bins):
    w = ro.window_table(bins)
    assert np.array_equal(w, R.nuttall_window(bins))
    assert w.max() <= 1.0                      # CPPAPP_ASSERT at src/FFTBackend.cpp:185
    if bins == 16384:                          # SURVEY.md 8(c) spot values
        assert w[0] == np.float32(-1.8626451e-09) and w[bins // 2] == 1.0
        assert abs(w.sum(dtype=np.float64) - 5828.5471) < 1e-3
        assert abs((w.astype(np.float64) ** 2).sum() - 4191.2492) < 1e-3


@pytest.mark.parametrize("n", [8, 1024, 65536, 524288])
def test_twiddle_table_accuracy(n):
    t = ro.twiddle_table(n).astype(np.complex128)
    j = np.arange(n)
    ref = np.exp(-2j * np.pi * j / n)
    assert np.abs(t - ref).max() < 6.1e-8      # half a float32 ulp of 1.0
    # exact symmetries of a correctly reduced table
    assert t[0] == 1.0 and t[n // 4] == -1j and t[n // 2] == -1.0
    assert np.array_equal(t[n // 2:], -t[:n // 2])


def test_last_error_reports_bad_arguments(lib):
    assert lib.ro_window_table(1, None) != 0
    assert b"ro_window_table" in lib.ro_last_error()
    assert lib.ro_version().startswith(b"radio_observer_b200")
