"""GPU parity of the chorus bank through the C ABI.  The reference has no chorus code -- only
the RNBO graph modules/rnbo/patcher/mono-chorus.rnbopat -- so the oracle (oracle/olb_oracle.cpp orc_chorus_*) is the
specification of the arithmetic: PARITY UNPINNED.  The bank must match it bit for bit.  The graph includes the gen~ pitch
shifter in front of the delay (pitchshift.gendsp, param pitch) and, for stereo-chorus.rnbopat, two mono choruses per instance."""
import numpy as np
import pytest
import torch

from ol_dsp_b200 import ChorusBank, StereoChorusBank

pytestmark = pytest.mark.gpu
NAMES = ("rate", "depth", "cutoff", "q", "mix", "phase", "pitch")


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def params_for(i, I):
    return dict(rate=(i % 11) / 10.0, depth=((i * 7) % 13) / 12.0, cutoff=((i * 3) % 17) / 16.0, q=((i * 5) % 9) / 8.0,
                mix=((i * 2) % 7) / 6.0, phase=(i % 5) / 4.0, pitch=[0.0, 3.0, 0.37, 1.0, 2.5, 0.01][i % 6])


def run_bank(x, plist, sr=48000.0, chunks=None, changes=None):
    T, I = x.shape
    b = ChorusBank(I, sample_rate=sr)
    for i, p in enumerate(plist):
        for k, v in (p or {}).items():
            b.Set(k, v, first=i, count=1)
    dx = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    dy = torch.empty_like(dx)
    a = 0
    for n in (chunks or [T]):
        for f, d in (changes or []):
            if f == a:
                for k, v in d.items():
                    b.Set(k, v)
        b.ProcessDevice(dx[a:].data_ptr(), dy[a:].data_ptr(), n)
        a += n
    assert a == T
    torch.cuda.synchronize()
    kern = b.LastKernel()
    b.close()
    return dy.cpu().numpy(), kern


@pytest.mark.parametrize("I,env", [(64, None), (84, None), (77, None), (64, "direct")])
def test_bank_matches_the_restatement(orc, monkeypatch, I, env):
    """Per-instance parameter sweep over the whole 0..1 range of every param (depth below its 0.08 floor included), ragged
    chunks whose sizes straddle the ring wrap (1 156 frames at 48 kHz), bank widths that are / are not a multiple of 4."""
    if env:
        monkeypatch.setenv("OLB_CHORUS_KERNEL", env)
    T = 6000
    x = orc.noise(21, 0, I, 0, T, 0.5)
    plist = [params_for(i, I) for i in range(I)]
    plist[1] = None  # patch defaults
    y, kern = run_bank(x, plist, chunks=[1, 63, 64, 1028, 1, 1155, 1157, T - 3469])
    assert kern == ("k_chorus_tma" if I % 4 == 0 and not env else "k_chorus_direct")
    for i in (0, 1, 2, 3, 10, 31, 32, 33, I - 1):
        want = orc.chorus(x[:, i], **(plist[i] or {}))
        assert np.array_equal(bits(y[:, i]), bits(want)), (I, i)


def test_parameter_change_mid_stream_reinit_and_in_place(orc):
    I, T = 8, 5000
    x = orc.noise(22, 0, I, 0, T, 0.5)
    ch = [(2000, dict(rate=0.9, depth=0.1, mix=1.0)), (3000, dict(cutoff=0.9, q=0.9, phase=0.25))]
    y, _ = run_bank(x, [None] * I, chunks=[2000, 1000, 2000], changes=ch)
    want = orc.chorus(x[:, 5], changes=ch)
    assert np.array_equal(bits(y[:, 5]), bits(want))
    b = ChorusBank(I, sample_rate=44100.0)
    dx = torch.from_numpy(x).cuda()
    b.ProcessDevice(dx.data_ptr(), dx.data_ptr(), 700)  # in place
    b.Init(48000.0)                                     # re-Init: ring, filter, phase cleared, another ring length
    dx = torch.from_numpy(x).cuda()
    b.ProcessDevice(dx.data_ptr(), dx.data_ptr(), T)
    torch.cuda.synchronize()
    assert np.array_equal(bits(dx[:, 3].cpu().numpy()), bits(orc.chorus(x[:, 3])))
    b.close()


def test_other_sample_rates_on_both_sides_of_the_shared_memory_limit(orc):
    """Up to ~53 kHz the ring is shared-memory resident for the launch; above (96 / 192 kHz: 2 324 / 4 628 frames x 32 lanes)
    the same kernel uses it in place in HBM.  Chunked, so the ring is carried from launch to launch in both forms."""
    x = orc.noise(23, 0, 40, 0, 6000, 0.5)
    for sr in (44100.0, 32000.0, 96000.0, 192000.0):
        y, _ = run_bank(x, [None] * 40, sr=sr, chunks=[1, 2047, 1500, 2452])
        for i in (2, 31, 32, 39):
            assert np.array_equal(bits(y[:, i]), bits(orc.chorus(x[:, i], sr=sr))), (sr, i)
    from ol_dsp_b200 import OlbError
    with pytest.raises(OlbError):
        ChorusBank(4, sample_rate=0.0)


def test_full_width_properties():
    """8 192 instances (BASELINE configs[4]): silence in -> silence out; mix = 0 returns the input bit for bit; identical
    instances give identical outputs."""
    I, T = 8192, 4096
    b = ChorusBank(I)
    z = torch.zeros((T, I), dtype=torch.float32, device="cuda")
    y = torch.empty_like(z)
    b.ProcessDevice(z.data_ptr(), y.data_ptr(), T)
    torch.cuda.synchronize()
    assert float(y.abs().max()) == 0.0
    b.Init(48000.0)
    b.Set("mix", 0.0)
    x = torch.rand((T, 1), device="cuda").expand(T, I).contiguous() - 0.5
    b.ProcessDevice(x.data_ptr(), y.data_ptr(), T)
    torch.cuda.synchronize()
    assert torch.equal(y, x)
    b.Init(48000.0)
    b.Set("mix", 0.5)
    b.ProcessDevice(x.data_ptr(), y.data_ptr(), T)
    torch.cuda.synchronize()
    assert torch.isfinite(y).all() and torch.equal(y, y[:, :1].expand(-1, I)) and not torch.equal(y, x)
    b.close()


@pytest.mark.parametrize("cut", [None, "100", "37", "640"])
def test_pitch_shifter_taps_reach_back_across_cuts_and_calls(orc, monkeypatch, cut):
    """The gen~ stage's taps reach up to one window (480 frames at 48 kHz) back: through the bank's input history when the
    block or the library's internal time cut is shorter than that.  Cuts of 37 and 100 frames, calls of 1 .. 2 000 frames,
    every instance with its own shifter frequency; in place."""
    if cut:
        monkeypatch.setenv("OLB_CHORUS_CUT", cut)
    I, T = 40, 5200
    x = orc.noise(25, 0, I, 0, T, 0.5)
    pitches = np.linspace(0.0, 3.0, I).astype(np.float32)
    b = ChorusBank(I)
    b.Set("pitch", pitches)
    b.Set("mix", 0.8)
    dx = torch.from_numpy(x.copy()).cuda()
    a = 0
    for n in (1, 2, 61, 300, 500, 2000, 479, 481, T - 3824):
        b.ProcessDevice(dx[a:].data_ptr(), dx[a:].data_ptr(), n)
        a += n
    assert a == T
    torch.cuda.synchronize()
    y = dx.cpu().numpy()
    b.close()
    for i in (0, 1, 13, 20, 39):
        want = orc.chorus(x[:, i], mix=0.8, pitch=float(pitches[i]))
        assert np.array_equal(bits(y[:, i]), bits(want)), (cut, i)
    # the stage is really in the path: at pitch 0 the wet signal is the input 240 frames late (+ the LFO delay)
    assert not np.array_equal(bits(y[:, 0]), bits(orc.chorus(x[:, 0], mix=0.8, pitch_stage=False)))


@pytest.mark.parametrize("sr", [44100.0, 96000.0])
def test_pitch_shifter_other_sample_rates(orc, sr):
    I, T = 8, 4000
    x = orc.noise(26, 0, I, 0, T, 0.5)
    y, _ = run_bank(x, [dict(pitch=1.5, rate=0.7)] * I, sr=sr, chunks=[1500, 2500])
    assert np.array_equal(bits(y[:, 6]), bits(orc.chorus(x[:, 6], sr=sr, pitch=1.5, rate=0.7)))


def test_pitch_shifter_resting_and_running_warps_in_one_bank(orc):
    """Warps whose 32 phasors all rest take the hoisted path of k_pitch_taps, the others the general one; a resting phasor
    that was running before (pitch set back to 0 mid-stream) rests at a phase other than 0."""
    I, T = 96, 3000
    x = orc.noise(28, 0, I, 0, T, 0.5)
    b = ChorusBank(I)
    b.Set("pitch", 2.0, first=32, count=33)  # warp 1 entirely, and lane 0 of warp 2
    dx = torch.from_numpy(x).cuda()
    dy = torch.empty_like(dx)
    b.ProcessDevice(dx.data_ptr(), dy.data_ptr(), 1700)
    b.Set("pitch", 0.0)                      # every phasor rests from here on, warp 1's at whatever phase it reached
    b.ProcessDevice(dx[1700:].data_ptr(), dy[1700:].data_ptr(), T - 1700)
    torch.cuda.synchronize()
    y = dy.cpu().numpy()
    b.close()
    for i, p in ((0, 0.0), (31, 0.0), (32, 2.0), (50, 2.0), (64, 2.0), (65, 0.0), (95, 0.0)):
        want = orc.chorus(x[:, i], pitch=p, changes=[(1700, dict(pitch=0.0))])
        assert np.array_equal(bits(y[:, i]), bits(want)), i


def test_stereo_chorus_is_two_mono_choruses_phase_1_and_0(orc):
    """stereo-chorus.rnbopat: in~ 1 -> mono-chorus (phase 1) -> out~ 1, in~ 2 -> mono-chorus (phase 0) -> out~ 2, the other
    params shared per stereo instance."""
    I, T = 18, 4000
    x = orc.noise(27, 0, 2 * I, 0, T, 0.5).reshape(T, I, 2)
    b = StereoChorusBank(I)
    rates = np.linspace(0.0, 1.0, I).astype(np.float32)
    b.Set("rate", rates)
    b.Set("depth", 0.9)
    b.Set("pitch", 0.5, first=3, count=2)
    y = b.ProcessHost(x)
    b.close()
    for i in (0, 3, 4, 17):
        pitch = 0.5 if i in (3, 4) else 0.0
        left = orc.chorus(x[:, i, 0], rate=float(rates[i]), depth=0.9, phase=1.0, pitch=pitch)
        right = orc.chorus(x[:, i, 1], rate=float(rates[i]), depth=0.9, phase=0.0, pitch=pitch)
        assert np.array_equal(bits(y[:, i, 0]), bits(left)) and np.array_equal(bits(y[:, i, 1]), bits(right)), i
