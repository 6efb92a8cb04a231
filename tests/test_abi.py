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


def test_cpp_host_mirror_builds_and_fails_loudly_without_gpu(tmp_path):
    """include/ol_gpu/BiquadBank.hpp: a reference-style C++ caller links against the C ABI; without a GPU its
    Init() reports the error instead of computing anything on the CPU."""
    import os
    import subprocess

    import __graft_entry__ as g
    g.build_cpp_hosts()
    exe = os.path.join(g.ROOT, "tests", "cpp", "biquad_host")
    assert os.path.exists(exe)
    if _lib.device_count() > 0:
        pytest.skip("a GPU is present")
    x = np.zeros((4, 8), dtype=np.float32)
    k = np.zeros((8, 8, 5), dtype=np.float32)
    x.tofile(tmp_path / "in.f32")
    k.tofile(tmp_path / "k.f32")
    r = subprocess.run([exe, "8", "4", str(tmp_path / "in.f32"), str(tmp_path / "k.f32"), str(tmp_path / "out.f32"), "0"],
                       capture_output=True, text=True)
    assert r.returncode == 4 and "no CPU fallback" in r.stderr
    # the fx / plate / chorus / voice mirrors (include/ol_gpu/{FxBank,DattorroBank,ChorusBank,VoiceBank}.hpp)
    exe = os.path.join(g.ROOT, "tests", "cpp", "units_host")
    assert os.path.exists(exe)
    x.tofile(tmp_path / "in.f32")
    for unit in ("rack", "plate", "chorus", "poly"):
        r = subprocess.run([exe, unit, "8", "4", str(tmp_path / "in.f32"), str(tmp_path / "out.f32"), str(tmp_path / "out2.f32")],
                           capture_output=True, text=True)
        assert r.returncode == 4 and "no CPU fallback" in r.stderr, (unit, r.returncode, r.stderr)


def test_product_never_touches_the_oracle():
    """Nothing under ol_dsp_b200/ or include/ may import, link or load oracle/ (the oracle is test infrastructure)."""
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    bad = []
    for top in ("ol_dsp_b200", "include"):
        for d, _, files in os.walk(os.path.join(root, top)):
            if "build" in d.split(os.sep) or "__pycache__" in d:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", "Makefile")):
                    txt = open(os.path.join(d, f), errors="ignore").read()
                    if re.search(r"(import|from)\s+oracle|oracle/|libolb_oracle|libolref", txt):
                        # comments that merely CITE the oracle file for op order are allowed
                        lines = [ln for ln in txt.splitlines() if re.search(r"oracle", ln) and not ln.lstrip().startswith(("//", "#", "*", "/*"))
                                 and "//" not in ln.split("oracle")[0]]
                        if lines:
                            bad.append((f, lines[:2]))
    assert not bad, bad
