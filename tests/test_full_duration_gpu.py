"""Parity over the FULL duration of the BASELINE configs (48 kHz x 10 s = 480 000 frames per unit), streamed in the block
sizes the benches use, for a handful of units of every bank against the oracle: bit-exact for the sequential-order
kernels, <= 1e-5 of full scale for the voice output (north_star's tolerance, the only device transcendental path).  The
unit counts are small so that the scalar oracle finishes in seconds; bank-width properties live in the per-unit files."""
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


@pytest.mark.parametrize("C,kernel,block", [(32768, "k_bq_tma", BLOCK), (8192, "k_bq_skew", BLOCK), (4096, "k_bq_skew", BLOCK), (4096, "k_bq_sys", 2400)])
def test_wide_banks_10s_fast(orc, C, kernel, block):
    """The throughput tier over the full 10 s (20 blocks of 24 000 frames, or 200 of 2 400) at bank widths and block lengths
    that reach each of the three TMA kernels through the automatic choice (k_bq_skew with its rolled tile in 96- and
    192-frame tiles at 8 192 and 4 096 channels); 64 channels spread over the bank against the oracle, bit-exact."""
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
