"""the reference's control wire format (modules/iolib/Serializer.h, modules/ctllib/Control.h)
restated in include/ol_gpu/ControlWire.hpp, pinned against the reference's OWN classes (oracle/_ref) where that library is
present and against committed golden records otherwise.  Host-only: no GPU."""
import os
import struct
import subprocess

import numpy as np
import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "control_wire.npz")
REC = 7 * 4 + 40


def tool():
    import __graft_entry__ as g
    g.build_cpp_hosts()
    return os.path.join(g.ROOT, "tests", "cpp", "wire_host")


def records(tmp_path):
    r = subprocess.run([tool(), "convert", str(tmp_path / "c.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    raw = open(tmp_path / "c.bin", "rb").read()
    assert len(raw) % REC == 0
    out = []
    for o in range(0, len(raw), REC):
        which, in_bits, ctl, adc, midi, fb, n = struct.unpack("<7I", raw[o:o + 28])
        out.append((which, in_bits, ctl, adc, midi, fb, raw[o + 28:o + 28 + n]))
    return out


def test_conversions_and_frames_match_the_reference_classes(ref, tmp_path):
    recs = records(tmp_path)
    assert len(recs) == 4097 + 128 + 1001
    if ref is not None and os.path.isdir("/root/reference"):
        for which, in_bits, ctl, adc, midi, fb, wire in recs:
            value = struct.unpack("<f", struct.pack("<I", in_bits))[0] if which == 1 else in_bits
            want = ref.control_wire(which, ctl, value)
            assert (ctl, adc, midi, fb, wire) == want, (which, in_bits)
    g = np.load(G)
    flat = np.array([[w, i, c, a, m, f] for w, i, c, a, m, f, _ in recs], dtype=np.uint32)
    assert np.array_equal(flat, g["records"])
    assert all(len(w) == 29 and w[:5] == b"\x2a" * 5 and struct.unpack("<3q", w[5:]) == (24, c, a) for _, _, c, a, _, _, w in recs)


def test_parser_reassembles_frames_from_ragged_pieces(tmp_path):
    recs = records(tmp_path)
    pick = recs[::97]
    stream = b"\x00\x07" + b"".join(w for *_, w in pick[:10]) + b"\x99" + b"".join(w for *_, w in pick[10:])
    stream += b"\x2a" * 6 + pick[0][6][5:]   # six start bytes: not a frame (the reference parses after exactly five)
    stream += pick[1][6]
    open(tmp_path / "in.bin", "wb").write(stream)
    r = subprocess.run([tool(), "parse", str(tmp_path / "in.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = np.fromfile(tmp_path / "out.bin", dtype=np.uint32).reshape(-1, 4)
    from_adc = {(c, a): (m, f) for w, _, c, a, m, f, _ in recs if w == 2}
    want = [(c, a) for _, _, c, a, _, _, _ in pick] + [(pick[1][2], pick[1][3])]
    assert [(int(c), int(a)) for c, a, _, _ in got] == want
    for c, a, m, f in got:  # a parsed frame carries the ADC value; MIDI and float views are derived from it
        if (int(c), int(a)) in from_adc:
            assert (int(m), int(f)) == from_adc[(int(c), int(a))]
