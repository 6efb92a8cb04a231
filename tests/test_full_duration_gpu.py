"""Parity over the FULL duration of the BASELINE configs (48 kHz x 10 s = 480 000 frames per unit), streamed in the block
sizes the benches use, against the oracle: bit-exact for the sequential-order kernels, <= 1e-5 of full scale for the voice
output (north_star's tolerance, the only device transcendental path).  First with a handful of units (every unit checked),
then with every bank at its BASELINE width (C3's 65 536 channels, C4's 65 536 voices, C5's 8 192 instances) on device-generated
noise, sampled units checked -- the scalar oracle runs only the sampled units, so the file still takes about a minute."""
import numpy as np
import pytest
import torch

from oracle import loader
from ol_dsp_b200 import MODE_EXACT, MODE_FAST, BiquadBank, ChorusBank, DattorroBank, FxBank, VoiceBank, c2_coefs, fx, noise_device

pytestmark = pytest.mark.gpu
T, BLOCK = 480000, 24000
NAN = float("nan")


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def stream(process, x, outs=1):
    """x [T][units] on the host -> the bank, BLOCK frames at a time (state stays on the device between blocks)."""
    dx = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    dys = [torch.empty_like(dx) for _ in range(outs)]
    for a in range(0, x.shape[0], BLOCK):
        n = min(BLOCK, x.shape[0] - a)
        process(dx[a:].data_ptr(), [d[a:].data_ptr() for d in dys], n)
    torch.cuda.synchronize()
    return [d.cpu().numpy() for d in dys]


@pytest.mark.parametrize("mode", [MODE_EXACT, MODE_FAST])
def test_biquad_cascade_10s(orc, mode):
    C = 64  # every coefficient set of the C2 / C3 workload (fc and gains cycle with channel mod 64)
    x = orc.noise(0x01D5B200, 0, C, 0, T)
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(mode)
    b.SetCoeffs(k)
    (y,) = stream(lambda i, o, n: b.ProcessDevice(i, o[0], n), x)
    b.close()
    want = orc.biquad(x, k, mode=mode)
    assert np.array_equal(bits(y), bits(want))
    if mode == MODE_FAST:  # the throughput tier against the bit-exact tier over the whole 10 s
        assert np.abs(y - orc.biquad(x, k, mode=MODE_EXACT)).max() <= 1e-3


def _streamed_bank_vs_oracle(orc, C, mode, kernel, blocks, block_frames, pick, seed=0x01D5B200):
    """A full-width bank fed device-generated Philox noise block by block (state resident between blocks); the channels in
    `pick` are checked against the oracle replaying the same noise.  Returns (max |y|, kernel name)."""
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(mode)
    b.SetCoeffs(k)
    dev = torch.empty((block_frames, C), dtype=torch.float32, device="cuda")
    out = torch.empty_like(dev)
    idx = torch.as_tensor(pick, device="cuda")
    st = np.zeros((len(pick), 8, 4), dtype=np.float32)
    kp = np.ascontiguousarray(k[pick])
    for blk in range(blocks):
        noise_device(dev.data_ptr(), seed, 0, C, blk * block_frames, block_frames)
        b.ProcessDevice(dev.data_ptr(), out.data_ptr(), block_frames)
        got = out[:, idx].cpu().numpy()
        x = np.stack([orc.noise(seed, c, 1, blk * block_frames, block_frames)[:, 0] for c in pick], axis=1)
        want = orc.biquad(x, kp, mode=mode, state=st)
        assert b.LastKernel() == kernel
        assert np.array_equal(bits(got), bits(want)), (blk, b.LastKernel())
    b.close()


@pytest.mark.parametrize("mode", [MODE_FAST, MODE_EXACT])
def test_headline_shape_one_block(orc, mode):
    """The shape bench.py times -- 65 536 channels x 24 000 frames through k_bq_tma -- with 256 channels spread over the
    bank (first, middle and last warps of the grid) checked bit for bit against the oracle."""
    pick = sorted(set(list(range(0, 64)) + list(range(20000, 20064)) + list(range(32768 + 31, 65536, 1021))[:64] + list(range(65472, 65536))))
    _streamed_bank_vs_oracle(orc, 65536, mode, "k_bq_tma", 1, 24000, pick)


@pytest.mark.parametrize("C,kernel,block", [(32768, "k_bq_tma", BLOCK), (16384, "k_bq_roll", BLOCK), (12288, "k_bq_skew", BLOCK), (8192, "k_bq_roll", BLOCK), (4096, "k_bq_skew", BLOCK),
                                               (4096, "k_bq_sys", 2400)])
def test_wide_banks_10s_fast(orc, C, kernel, block):
    """The throughput tier over the full 10 s (20 blocks of 24 000 frames, or 200 of 2 400) at bank widths and block lengths
    that reach each of the four TMA kernels through the automatic choice (k_bq_roll with two and four sections per warp at
    8 192 and 16 384 channels, k_bq_skew with its unrolled and rolled tiles at 12 288 and 4 096); 64 channels spread over the bank against the oracle, bit-exact."""
    pick = sorted(set(list(range(0, 16)) + list(range(C // 2 - 8, C // 2 + 8)) + list(range(7, C, C // 16))[:16] + list(range(C - 16, C))))
    _streamed_bank_vs_oracle(orc, C, MODE_FAST, kernel, T // block, block, pick)


@pytest.mark.parametrize("kind,name", [(fx.DELAY, "delayfx"), (fx.RACK, "fxrack")])
def test_fx_units_10s(orc, kind, name):
    I = 4
    x = orc.noise(0x01D5B200, 0, I, 0, T, 0.25)
    times = [0.05, 0.35, 0.65, 0.95]  # the C5 delay-time sweep
    b = FxBank(kind, I)
    b.UpdateHardwareControl(35, np.asarray(times, dtype=np.float32))
    (y,) = stream(lambda i, o, n: b.ProcessDevice(i, o[0], n), x)
    b.close()
    for i in range(I):
        want = getattr(orc, name)(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, [(0, loader.EV_HW_CC, 35, times[i])])
        assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), i


def test_dattorro_10s(orc):
    I = 4
    x = orc.noise(0x01D5B200, 0, I, 0, T, 0.25)
    decay = [0.5, 0.6, 0.75, 0.875]  # the C5 decay sweep
    b = DattorroBank(I)
    b.SetDecay(np.asarray(decay, dtype=np.float32))
    l, r = stream(lambda i, o, n: b.ProcessDevice(i, o[0], o[1], n), x, outs=2)
    b.close()
    for i in range(I):
        wl, wr = orc.dattorro(x[:, i], [NAN, NAN, NAN, NAN, decay[i], NAN, NAN])
        assert np.array_equal(bits(l[:, i]), bits(wl)) and np.array_equal(bits(r[:, i]), bits(wr)), i
    assert np.isfinite(l).all() and np.abs(l[-48000:]).max() > 0  # the 16-bit counter wrapped 7 times, the tank still rings


def test_chorus_10s(orc):
    I = 4
    x = orc.noise(0x01D5B200, 0, I, 0, T, 0.25)
    p = [dict(rate=0.0, depth=0.08), dict(rate=0.3, depth=0.5), dict(rate=0.7, depth=0.9), dict(rate=1.0, depth=1.0, phase=0.5)]
    b = ChorusBank(I)
    for i, d in enumerate(p):
        for k, v in d.items():
            b.Set(k, v, first=i, count=1)
    (y,) = stream(lambda i, o, n: b.ProcessDevice(i, o[0], n), x)
    b.close()
    for i in range(I):
        assert np.array_equal(bits(y[:, i]), bits(orc.chorus(x[:, i], **p[i]))), i


def test_voices_10s(orc):
    """C4's script: NoteOn at t = 0, NoteOff at 5.0 s, presets through the MIDI maps; notes 36, 60, 84, 96."""
    notes = [36, 60, 84, 96]
    V = len(notes)
    ccs = [(41, 70), (42, 30), (73, 64), (75, 60), (108, 10), (111, 40)]
    b = VoiceBank(V)
    for cc, val in ccs:
        b.UpdateMidiControl(cc, val)
    b.NoteOn(np.asarray(notes, dtype=np.uint8))
    out = torch.empty((T, V), dtype=torch.float32, device="cuda")
    for a in range(0, T, BLOCK):
        if a == T // 2:
            b.NoteOff()
        b.ProcessDevice(BLOCK, out_ptr=out[a:].data_ptr())
    torch.cuda.synchronize()
    b.close()
    y = out.cpu().numpy()
    for i, note in enumerate(notes):
        ev = [(0, loader.EV_MIDI_CC, cc, val) for cc, val in ccs] + [(0, loader.EV_NOTE_ON, note, 100), (T // 2, loader.EV_NOTE_OFF, note, 0)]
        want = orc.voice(T, ev=ev)
        assert np.abs(y[:, i] - want).max() <= 1e-5, (note, float(np.abs(y[:, i] - want).max()))
        assert np.abs(want[:T // 2]).max() > 0.01  # the voice sounded


# ---- banks at their BASELINE width over long runs: device-generated noise block by block, sampled units against the oracle ----
def _wide_stream(process, units, frames, block, pick, scale=0.25, seed=0x01D5B200, outs=1):
    """Streams `frames` frames of Philox noise through a bank of `units` units, `block` frames at a time, keeping only the
    columns in `pick`.  Returns (x[frames][len(pick)], [y ...]) on the host."""
    dev = torch.empty((block, units), dtype=torch.float32, device="cuda")
    dout = [torch.empty_like(dev) for _ in range(outs)]
    idx = torch.as_tensor(pick, device="cuda")
    xs, ys = [], [[] for _ in range(outs)]
    for a in range(0, frames, block):
        n = min(block, frames - a)
        noise_device(dev.data_ptr(), seed, 0, units, a, n, scale)
        process(dev.data_ptr(), [d.data_ptr() for d in dout], n)
        xs.append(dev[:n, idx].cpu().numpy())
        for k in range(outs):
            ys[k].append(dout[k][:n, idx].cpu().numpy())
    return np.concatenate(xs), [np.concatenate(y) for y in ys]


@pytest.mark.parametrize("kind,name,I,frames,kernel", [(fx.RACK, "fxrack", 8192, T, "k_fx_split"), (fx.DELAY, "delayfx", 8192, T, "k_fx_split"),
                                                       (fx.RACK, "fxrack", 16384, T, "k_fx_split"), (fx.FILTER, "filterfx", 65536, T, "k_fx_tma")])
def test_fx_banks_at_baseline_width_long(orc, kind, name, I, frames, kernel):
    """C5's 8 192 instances (FxRack, DelayFx), 16 384 racks (four pipelines per CTA) and 65 536 filters over the full 10 s:
    the write pointer of the 48 000-sample delay lines wraps 10 times under the warp pipelines with 2 and 4 pipelines per CTA.
    C5's delay-time sweep across the bank; sampled instances bit-exact against the oracle."""
    b = FxBank(kind, I)
    times = (0.05 + 0.9 * np.arange(I) / I).astype(np.float32)
    b.UpdateHardwareControl(35, times)  # CC_DELAY_TIME
    pick = sorted({0, 1, 31, 32, I // 2 - 1, I // 2, I - 33, I - 1})
    x, (y,) = _wide_stream(lambda i, o, n: b.ProcessDevice(i, o[0], n), I, frames, BLOCK, pick)
    assert b.LastKernel() == kernel
    b.close()
    for j, i in enumerate(pick):
        want = getattr(orc, name)(np.ascontiguousarray(x[:, j:j + 1]), 48000.0, [(0, loader.EV_HW_CC, 35, float(times[i]))])
        assert np.array_equal(bits(y[:, j]), bits(want[:, 0])), (name, I, i)


def test_dattorro_at_baseline_width_across_the_counter_wrap(orc):
    """C5's 8 192 plates (two waves of the four-warp kernel with its rings in HBM) over the full 10 s: the 16-bit sample counter
    all ring indices derive from wraps 7 times.  C5's decay sweep; sampled instances bit-exact against verb.cpp's restatement."""
    I, frames = 8192, T
    b = DattorroBank(I)
    decay = (0.5 + 0.4 * (np.arange(I) % 16) / 16).astype(np.float32)
    b.SetDecay(decay)
    pick = sorted({0, 31, 32, 4735, 4736, I - 1})
    x, (l, r) = _wide_stream(lambda i, o, n: b.ProcessDevice(i, o[0], o[1], n), I, frames, BLOCK, pick, outs=2)
    assert b.LastKernel() == "k_dattorro_split"
    b.close()
    for j, i in enumerate(pick):
        wl, wr = orc.dattorro(x[:, j], [NAN, NAN, NAN, NAN, float(decay[i]), NAN, NAN])
        assert np.array_equal(bits(l[:, j]), bits(wl)) and np.array_equal(bits(r[:, j]), bits(wr)), i


def test_chorus_at_baseline_width_long(orc):
    """C5's 8 192 choruses over the full 10 s, rate and depth swept across the bank, every 7th instance with a running pitch shifter."""
    I, frames = 8192, T
    b = ChorusBank(I)
    rate = (np.arange(I) % 101 / 100.0).astype(np.float32)
    depth = (0.08 + 0.92 * (np.arange(I) % 53) / 52.0).astype(np.float32)
    pitch = np.where(np.arange(I) % 7 == 0, 0.25 + (np.arange(I) % 11) / 4.0, 0.0).astype(np.float32)
    b.Set("rate", rate)
    b.Set("depth", depth)
    b.Set("pitch", pitch)
    pick = sorted({0, 7, 31, 32, 4095, 4102, I - 2, I - 1})
    x, (y,) = _wide_stream(lambda i, o, n: b.ProcessDevice(i, o[0], n), I, frames, BLOCK, pick)
    b.close()
    for j, i in enumerate(pick):
        want = orc.chorus(x[:, j], rate=float(rate[i]), depth=float(depth[i]), pitch=float(pitch[i]))
        assert np.array_equal(bits(y[:, j]), bits(want)), i


def test_voices_at_baseline_width(orc):
    """C4's 65 536 voices (note 36 + v mod 61) over the full 10 s in 4 800-frame blocks, NoteOff at 5 s; sampled voices (first,
    last, both sides of the note wrap and of the bank's middle) within 1e-5 of the oracle."""
    V, frames, block = 65536, T, 4800
    notes = (36 + np.arange(V) % 61).astype(np.uint8)
    ccs = [(41, 70), (42, 30), (73, 64), (75, 60), (108, 10), (111, 40)]
    b = VoiceBank(V)
    for cc, val in ccs:
        b.UpdateMidiControl(cc, val)
    b.NoteOn(notes)
    pick = sorted({0, 1, 60, 61, 32767, 32768, V - 2, V - 1})
    idx = torch.as_tensor(pick, device="cuda")
    out = torch.empty((block, V), dtype=torch.float32, device="cuda")
    ys = []
    for a in range(0, frames, block):
        if a == frames // 2:
            b.NoteOff()
        b.ProcessDevice(block, out_ptr=out.data_ptr())
        ys.append(out[:, idx].cpu().numpy())
    b.close()
    y = np.concatenate(ys)
    for j, v in enumerate(pick):
        note = int(notes[v])
        ev = [(0, loader.EV_MIDI_CC, cc, val) for cc, val in ccs] + [(0, loader.EV_NOTE_ON, note, 100), (frames // 2, loader.EV_NOTE_OFF, note, 0)]
        want = orc.voice(frames, ev=ev)
        assert np.abs(y[:, j] - want).max() <= 1e-5, (v, note, float(np.abs(y[:, j] - want).max()))
