"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol include/olbank.h
declares, validates arguments, and fails loudly (no fallback) without a GPU."""
import ctypes as C

import numpy as np
import pytest

from ol_dsp_b200 import _lib


def test_every_declared_symbol_is_exported():
    L = _lib.lib()
    names = _lib.declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/olbank.h but not exported by libolbank.so"


def test_version_and_counters():
    L = _lib.lib()
    assert b"sm_100a" in L.olb_version()
    assert _lib.launch_count() >= 0


def test_design_matches_oracle(orc):
    from ol_dsp_b200 import design
    for typ, f0, q, g in [(0, 1000.0, 0.707, 0.0), (1, 30.0, 0.707, 0.0), (2, 2500.0, 2.0, 3.0), (2, 400.0, 2.0, -3.0), (3, 5000.0, 1.0, 0.0)]:
        a = design(typ, 48000.0, f0, q, g)
        b = orc.rbj(typ, 48000.0, f0, q, g)
        assert a.tobytes() == b.tobytes()


def test_bad_arguments_are_rejected_not_crashing():
    L = _lib.lib()
    h = C.c_void_p()
    assert L.olb_biquad_create(0, 0, 8, 48000.0, C.byref(h)) == _lib.E_ARG
    assert L.olb_biquad_create(0, 16, 0, 48000.0, C.byref(h)) == _lib.E_ARG
    assert L.olb_biquad_create(0, 16, 99, 48000.0, C.byref(h)) == _lib.E_ARG
    assert L.olb_biquad_process(None, None, None, 1, 0, None) == _lib.E_ARG
    assert b"NULL" in L.olb_last_error()
    out = np.zeros(5, dtype=np.float32)
    assert L.olb_biquad_design(7, 48000.0, 100.0, 1.0, 0.0, out.ctypes.data_as(C.POINTER(C.c_float))) == _lib.E_ARG


def test_no_cpu_fallback_without_gpu():
    if _lib.device_count() > 0:
        pytest.skip("a GPU is present")
    from ol_dsp_b200 import BiquadBank, OlbError
    with pytest.raises(OlbError) as e:
        BiquadBank(64, 8)
    assert e.value.code == _lib.E_CUDA
    assert "no CPU fallback" in str(e.value)
