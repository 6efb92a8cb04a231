"""C++ callers written against include/ol_gpu/*.hpp (the reference-style host mirrors: Init -> setters -> Process per frame
-> ProcessBlock, host buffers) run on the GPU box and compared with the oracle."""
import os
import subprocess

import numpy as np
import pytest

from oracle import loader

pytestmark = pytest.mark.gpu
NAN = float("nan")


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def run(unit, I, T, x, tmp_path, two=False):
    import __graft_entry__ as g
    g.build_cpp_hosts()
    exe = os.path.join(g.ROOT, "tests", "cpp", "units_host")
    if x is not None:
        x.tofile(tmp_path / "in.f32")
    args = [exe, unit, str(I), str(T), str(tmp_path / "in.f32"), str(tmp_path / "out.f32")] + ([str(tmp_path / "out2.f32")] if two else [])
    r = subprocess.run(args, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return r.stdout


@pytest.mark.parametrize("unit,name", [("filter", "filterfx"), ("delay", "delayfx"), ("reverb", "reverbfx"), ("rack", "fxrack")])
def test_fx_mirrors(orc, tmp_path, unit, name):
    I, T = 40, 700
    x = orc.noise(51, 0, I, 0, T, 0.5)
    run(unit, I, T, x, tmp_path)
    y = np.fromfile(tmp_path / "out.f32", dtype=np.float32).reshape(T, I)
    ev = [(0, loader.EV_HW_CC, 35, 0.3), (0, loader.EV_MIDI_CC, 41, 90)]
    for i in (0, 7, 39):
        want = getattr(orc, name)(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, ev)
        assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), (unit, i)


def test_plate_mirror(orc, tmp_path):
    I, T = 12, 900
    x = orc.noise(52, 0, I, 0, T, 0.3)
    run("plate", I, T, x, tmp_path, two=True)
    l = np.fromfile(tmp_path / "out.f32", dtype=np.float32).reshape(T, I)
    r = np.fromfile(tmp_path / "out2.f32", dtype=np.float32).reshape(T, I)
    wl, wr = orc.dattorro(x[:, 5], [0.05, NAN, NAN, NAN, 0.8, NAN, NAN])
    assert np.array_equal(bits(l[:, 5]), bits(wl)) and np.array_equal(bits(r[:, 5]), bits(wr))


def test_chorus_mirror(orc, tmp_path):
    I, T = 16, 2000
    x = orc.noise(53, 0, I, 0, T, 0.5)
    run("chorus", I, T, x, tmp_path)
    y = np.fromfile(tmp_path / "out.f32", dtype=np.float32).reshape(T, I)
    assert np.array_equal(bits(y[:, 9]), bits(orc.chorus(x[:, 9], rate=0.7, depth=0.9)))


@pytest.mark.parametrize("unit,name", [("rack2", "fxrack"), ("delay2", "delayfx")])
def test_stereo_fx_mirrors(orc, tmp_path, unit, name):
    """ol::gpu::FxRack2Bank / DelayFx2Bank: the reference's <2> templates, frames [instances][2]."""
    I, T = 12, 900
    x = orc.noise(54, 0, 2 * I, 0, T, 0.5)
    run(unit, 2 * I, T, x, tmp_path)
    y = np.fromfile(tmp_path / "out.f32", dtype=np.float32).reshape(T, I, 2)
    ev = [(0, loader.EV_HW_CC, 35, 0.3), (0, loader.EV_MIDI_CC, 41, 90)]
    for i in (0, 5, 11):
        want = getattr(orc, name)(np.ascontiguousarray(x.reshape(T, I, 2)[:, i, :]), 48000.0, ev)
        assert np.array_equal(bits(y[:, i, :]), bits(want)), (unit, i)


def test_fx_mirrors_in_the_fused_tier(orc, tmp_path):
    """FxRackBank::SetMode(OLB_MODE_FMA) / FxRack2Bank::SetMode: the C++ mirrors reach olb_fx_set_mode / olb_fx2_set_mode."""
    I, T = 40, 700
    ev = [(0, loader.EV_HW_CC, 35, 0.3), (0, loader.EV_MIDI_CC, 41, 90)]
    x = orc.noise(55, 0, I, 0, T, 0.5)
    run("rack_fma", I, T, x, tmp_path)
    y = np.fromfile(tmp_path / "out.f32", dtype=np.float32).reshape(T, I)
    run("rack2_fma", I, T, x, tmp_path)
    y2 = np.fromfile(tmp_path / "out.f32", dtype=np.float32).reshape(T, I // 2, 2)
    exact = orc.fxrack(np.ascontiguousarray(x[:, 7:8]), 48000.0, ev)
    try:
        orc.fx_mode(True)
        for i in (0, 7, 39):
            want = orc.fxrack(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, ev)
            assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), i
        for i in (0, 5, 19):
            want = orc.fxrack(np.ascontiguousarray(x.reshape(T, I // 2, 2)[:, i, :]), 48000.0, ev)
            assert np.array_equal(bits(y2[:, i, :]), bits(want)), i
    finally:
        orc.fx_mode(False)
    assert not np.array_equal(bits(y[:, 7]), bits(exact[:, 0]))


def test_stereo_chorus_mirror(orc, tmp_path):
    """ol::gpu::StereoChorusBank (stereo-chorus.rnbopat): left = phase 1, right = phase 0, shared rate / depth / pitch."""
    I, T = 6, 2000
    x = orc.noise(55, 0, 2 * I, 0, T, 0.5)
    run("chorus2", 2 * I, T, x, tmp_path)
    y = np.fromfile(tmp_path / "out.f32", dtype=np.float32).reshape(T, I, 2)
    x3 = x.reshape(T, I, 2)
    for i in (0, 5):
        assert np.array_equal(bits(y[:, i, 0]), bits(orc.chorus(x3[:, i, 0], rate=0.7, depth=0.9, pitch=0.5, phase=1.0)))
        assert np.array_equal(bits(y[:, i, 1]), bits(orc.chorus(x3[:, i, 1], rate=0.7, depth=0.9, pitch=0.5, phase=0.0)))


def test_polyvoice_mirror(orc, tmp_path):
    V, T = 4, 3000
    out = run("poly", V, T, None, tmp_path)
    y = np.fromfile(tmp_path / "out.f32", dtype=np.float32)
    ev = [(0, loader.EV_NOTE_ON, 48, 100), (0, loader.EV_NOTE_ON, 55, 100), (T // 2, loader.EV_NOTE_OFF, 48, 0), (T // 2, loader.EV_NOTE_ON, 64, 100)]
    want = orc.polyvoice(V, T, ev=ev)
    assert np.abs(y - np.asarray(want).reshape(-1)).max() <= 4e-5
    assert out.split()[:4] == ["64", "55", "0", "0"]  # NoteOff(48) frees voice 0, NoteOn(64) takes the first free voice


def test_voicemap_mirror(tmp_path):
    """ol::gpu::VoiceMapBank driven like the reference's VoiceMap, against the reference's own output (golden)."""
    want = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "synth.npz"))["voicemap"]
    run("voicemap", 6, want.shape[0], None, tmp_path)
    y = np.fromfile(tmp_path / "out.f32", dtype=np.float32)
    assert np.abs(y - want).max() <= 5e-5


def test_sampler_mirror(orc, tmp_path):
    """ol::gpu::SamplerBank driven like the reference's Sample / SampleSoundSource, against the oracle restatement."""
    I, T, n = 40, 500, 300
    data = orc.noise(54, 0, 1, 0, n, 0.5)[:, 0].copy()
    out = run("sampler", I, T, data, tmp_path)
    y = np.fromfile(tmp_path / "out.f32", dtype=np.float32).reshape(T, I)
    L = loader
    for i in (0, 1, 7, 39):
        ev = [(0, L.EV_GATE_ON, 0, 0.0), (0, L.EV_SAMPLE_SEEK, n - 10, 0.0)] if i == 0 else \
             [(0, L.EV_SAMPLE_MODE, 1, 0.0), (0, L.EV_SAMPLE_LOOP_START, i, 0.0), (0, L.EV_SAMPLE_LOOP_END, i + 40, 0.0), (0, L.EV_GATE_ON, 0, 0.0)]
        want = np.full((T, 1), 9.0, dtype=np.float32)
        st = orc.sampler(data, ev, want)
        assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), i
        if i == 0:
            assert int(out.split()[0]) == int(st[4]) == n


def test_setters_from_a_control_thread_never_tear_a_block():
    """tests/cpp/setter_race.cpp: a control thread flips the coefficient table of every channel (and an fx bank's cutoff)
    as fast as it can while the audio thread streams blocks; every block must be computed with ONE coherent table."""
    import __graft_entry__ as g
    g.build_cpp_hosts()
    r = subprocess.run([os.path.join(g.ROOT, "tests", "cpp", "setter_race"), "4096", "300", "64"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "torn 0" in r.stdout
