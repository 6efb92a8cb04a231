"""SURVEY.md section 8f item 1 / BASELINE configs[0] on the GPU: a WAV file through the fx rack, end to end through the C++
host mirror (tests/cpp/wav_rack.cpp: include/ol_gpu/Wav.hpp + FxBank.hpp).  The reference's material is
workouts/drum_loop.wav (stereo, 24-bit PCM, 44.1 kHz); it cannot travel to the GPU box, so the tests synthesise a file of
that format.  The oracle for the plate rack is the composition of pieces that are each pinned on their own: DelayFx ->
Dattorro plate (left) in ReverbFx's dry/wet -> FilterFx -> master volume."""
import os
import struct
import subprocess

import numpy as np
import pytest

NAN = float("nan")


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def write_pcm24(path, ints, sr):
    frames, ch = ints.shape
    data = b"".join(struct.pack("<i", int(v))[:3] for v in ints.reshape(-1))
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVE" + b"fmt " + struct.pack("<IHHIIHH", 16, 1, ch, sr, sr * ch * 3, ch * 3, 24)
                + b"data" + struct.pack("<I", len(data)) + data)


def read_f32(path):
    raw = open(path, "rb").read()
    assert raw[:4] == b"RIFF" and raw[8:12] == b"WAVE" and raw[12:16] == b"fmt "
    fmt, ch, sr = struct.unpack("<HHI", raw[20:28])
    assert fmt == 3 and raw[36:40] == b"data"
    n = struct.unpack("<I", raw[40:44])[0]
    return np.frombuffer(raw[44:44 + n], dtype=np.float32).reshape(-1, ch).copy(), sr


def tool():
    import __graft_entry__ as g
    g.build_cpp_hosts()
    return os.path.join(g.ROOT, "tests", "cpp", "wav_rack")


def material(frames=3000, seed=5):
    rng = np.random.default_rng(seed)
    ints = rng.integers(-(1 << 22), 1 << 22, size=(frames, 2), dtype=np.int64)
    ints[0] = [-(1 << 23), (1 << 23) - 1]  # the extremes of 24-bit PCM
    return ints


def test_wav_reader_decodes_pcm24_like_miniaudio(tmp_path):
    """CPU: s24 -> f32 is x / 2^23 (what miniaudio's ma_pcm_s24_to_f32 computes), interleaving and sample rate kept."""
    ints = material()
    write_pcm24(tmp_path / "in.wav", ints, 44100)
    r = subprocess.run([tool(), str(tmp_path / "in.wav"), str(tmp_path / "out.wav"), "decode"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    y, sr = read_f32(tmp_path / "out.wav")
    assert sr == 44100 and np.array_equal(bits(y), bits((ints / 8388608.0).astype(np.float32)))


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["stub", "plate"])
def test_wav_through_the_rack(orc, tmp_path, mode):
    ints = material()
    write_pcm24(tmp_path / "in.wav", ints, 44100)
    r = subprocess.run([tool(), str(tmp_path / "in.wav"), str(tmp_path / "out.wav"), mode, "1000"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    y, sr = read_f32(tmp_path / "out.wav")
    x = (ints / 8388608.0).astype(np.float32)
    for c in range(2):
        xc = np.ascontiguousarray(x[:, c:c + 1])
        if mode == "stub":
            want = orc.fxrack(xc, 44100.0, [])[:, 0]
        else:
            d = orc.delayfx(xc, 44100.0, [])[:, 0]
            damping = orc.scale(12000.0, 0.0, 20000.0, 0.0, 1.0, 1.0)  # ReverbFx::cutoff default, back to 0..1
            wl, _ = orc.dattorro(d, [NAN, 0.5, 0.5, 0.5, 0.5, 0.5, damping])
            b = np.float32(0.1)  # ReverbFx::balance default
            mix = wl * b + d * (np.float32(1.0) - b)
            want = orc.filterfx(np.ascontiguousarray(mix.reshape(-1, 1)), 44100.0, [])[:, 0] * np.float32(0.8)
        assert np.array_equal(bits(y[:, c]), bits(want)), (mode, c)
