"""GPU parity of the fxlib unit banks through the C ABI: bit-exact against the oracle,
and against tests/golden/fx.npz -- outputs of the reference's OWN headers (modules/fxlib/Fx.h compiled unmodified over
the DaisySP restatement; Svf / DelayLine arithmetic is the unpinned part)."""
import hashlib
import os

import numpy as np
import pytest
import torch

from oracle import loader
from ol_dsp_b200 import MODE_EXACT, MODE_FMA, FxBank, _lib, fx
from tests.golden import make_golden as mg

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAMES = {fx.FILTER: "filterfx", fx.DELAY: "delayfx", fx.REVERB: "reverbfx", fx.RACK: "fxrack"}


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run_bank(kind, x, events_per_instance, sr=48000.0, extra_cuts=(), mode=None):
    """x: [frames][instances].  events_per_instance[i] = list of (frame, kind, cc, value) applied before that frame;
    the block is cut at every event frame (setters take effect at the next Process call)."""
    T, I = x.shape
    b = FxBank(kind, I, sample_rate=sr)
    if mode is not None:
        b.SetMode(mode)
    cuts = sorted({0, T} | {e[0] for ev in events_per_instance for e in ev} | set(extra_cuts))
    dx = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    dy = torch.empty_like(dx)
    for a, z in zip(cuts[:-1], cuts[1:]):
        for i, ev in enumerate(events_per_instance):
            for (f, k, cc, v) in ev:
                if f != a:
                    continue
                if k == loader.EV_MIDI_CC:
                    b.UpdateMidiControl(cc, int(v), first=i, count=1)
                elif k == loader.EV_HW_CC:
                    b.UpdateHardwareControl(cc, v, first=i, count=1)
                elif k == loader.EV_INIT:
                    b.Init(v)
        b.ProcessDevice(dx[a:].data_ptr(), dy[a:].data_ptr(), z - a)
    torch.cuda.synchronize()
    b.close()
    return dy.cpu().numpy()


@pytest.mark.parametrize("kind", [fx.FILTER, fx.DELAY, fx.REVERB, fx.RACK])
def test_unit_matches_reference_golden(orc, kind):
    """tests/golden/fx.npz: the reference's own class fed 16 000 frames of noise with the FX_EVENTS control script."""
    g = np.load(os.path.join(G, "fx.npz"))
    seed, frames, scale = g["meta"]
    x = orc.noise(int(seed), 0, 1, 0, int(frames), float(scale))
    xs = np.concatenate([x, orc.noise(5, 1, 2, 0, int(frames), 0.4)], axis=1)  # two bystander instances
    y = run_bank(kind, xs, [mg.FX_EVENTS, [], [(0, loader.EV_HW_CC, 35, 0.013)]], extra_cuts=(1, 777))
    name = NAMES[kind]
    assert np.array_equal(bits(y[:512, 0]), bits(g[f"{name}_head"]))
    assert np.array_equal(bits(y[-512:, 0]), bits(g[f"{name}_tail"]))
    assert sha(np.ascontiguousarray(y[:, :1])) == str(g[f"{name}_sha"][0])
    for i, ev in ((1, []), (2, [(0, loader.EV_HW_CC, 35, 0.013)])):
        want = getattr(orc, name)(np.ascontiguousarray(xs[:, i:i + 1]), 48000.0, ev)
        assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), (name, i)


def test_delay_bank_per_instance_times_filter_types_and_reinit(orc):
    """C5-style sweep: delay time differs per instance (including 0 and the maximum), wet filter type / drive vary,
    then a re-Init mid-stream (test/fx_test.cpp:32,42) -- all against the oracle."""
    I, T = 70, 6000
    x = orc.noise(9, 0, I, 0, T, 0.5)
    evs = []
    for i in range(I):
        t = [0.0, 1.0, 0.0001, 0.05 + 0.9 * i / I][min(i, 3)]
        ev = [(0, loader.EV_HW_CC, 35, t), (0, loader.EV_MIDI_CC, 36, 30 + i), (0, loader.EV_MIDI_CC, 37, 20 + i),
              (2000, loader.EV_MIDI_CC, 38, (i * 7) % 128)]
        if i % 5 == 0:
            ev.append((4000, loader.EV_INIT, 0, 44100.0))
        evs.append(ev)
    # Init is bank-wide in the C ABI: give every instance the same re-Init so the oracle comparison is per instance
    evs = [[e for e in ev if e[1] != loader.EV_INIT] + [(4000, loader.EV_INIT, 0, 44100.0)] for ev in evs]
    y = run_bank(fx.DELAY, x, evs)
    for i in (0, 1, 2, 3, 33, 69):
        want = orc.delayfx(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, evs[i])
        assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), i


def test_filter_bank_all_types_and_hw_controls(orc):
    I, T = 40, 3000
    x = orc.noise(10, 0, I, 0, T, 0.8)
    evs = [[(0, loader.EV_MIDI_CC, 43, (i * 26) % 128), (0, loader.EV_HW_CC, 41, 0.02 + 0.9 * i / I), (0, loader.EV_HW_CC, 42, (i % 10) / 10.0),
            (1000, loader.EV_HW_CC, 44, (i % 7) / 7.0)] for i in range(I)]
    y = run_bank(fx.FILTER, x, evs)
    for i in range(I):
        want = orc.filterfx(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, evs[i])
        assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), i


def test_reference_fx_test_delay_replayed_on_the_bank(orc):
    """test/fx_test.cpp:25-55 on the bank: Init(128), one Process, UpdateHardwareControl(CC_DELAY_TIME, 0.5),
    Init(48000), 48 000 frames of a 20 kHz sine -> no NaN (the reference's only assertion) and bit-equal to the oracle."""
    n = 48000
    ph = (np.arange(n, dtype=np.float64) * 20000.0 / 48000.0) % 1.0
    sine = (0.5 * np.sin(2 * np.pi * ph)).astype(np.float32).reshape(n, 1)
    b = FxBank(fx.DELAY, 1, sample_rate=128.0)
    one = torch.zeros((1, 1), dtype=torch.float32, device="cuda")
    b.ProcessDevice(one.data_ptr(), one.data_ptr(), 1)
    b.UpdateHardwareControl(35, 0.5)
    b.Init(48000.0)
    dx = torch.from_numpy(sine).cuda()
    b.ProcessDevice(dx.data_ptr(), dx.data_ptr(), n)  # in place, as main.cpp:82-85 calls it
    torch.cuda.synchronize()
    y = dx.cpu().numpy()
    assert not np.isnan(y).any()
    # the oracle run includes the first throw-away frame at index 0; replay it the same way
    full = np.concatenate([np.zeros((1, 1), np.float32), sine])
    want = orc.delayfx(full, 128.0, [(1, loader.EV_HW_CC, 35, 0.5), (1, loader.EV_INIT, 0, 48000.0)])[1:]
    assert np.array_equal(bits(y), bits(want))
    b.close()


VARIANTS = [{"OLB_FX_KERNEL": "simple"}, {"OLB_FX_KERNEL": "pf"}, {"OLB_FX_KERNEL": "tma"}] + \
    [{"OLB_FX_KERNEL": "tma", "OLB_FX_SPLIT": "0", "OLB_FX_TMA_CFG": c} for c in ("32x4", "32x8", "64x8")] + \
    [{"OLB_FX_KERNEL": "tma", "OLB_FX_SPLIT_G": g, "OLB_FX_TMA_CFG": c} for g, c in (("1", "32x4"), ("2", "64x4"), ("2", "32x4"), ("4", "32x4"))]
# FilterFx: "tma" is k_fx_tma in the shape OLB_FX_TMA_CFG names.  Ring kinds: "tma" is the warp pipeline k_fx_split with
# OLB_FX_SPLIT_G pipelines per CTA, unless OLB_FX_SPLIT=0 asks for the one-warp k_fx_tma.


@pytest.mark.parametrize("env", VARIANTS, ids=lambda e: "-".join(e.values()))
@pytest.mark.parametrize("kind", [fx.FILTER, fx.DELAY, fx.RACK])
def test_every_kernel_variant_is_bit_exact(orc, monkeypatch, env, kind):
    """The one-frame kernel, the register double-buffered kernel and the TMA-pipelined kernel (every compiled tile shape)
    compute the same bits: delays on both sides of the look-ahead thresholds (ring reads run ahead only past 2B + 2 resp.
    STAGES x TB + 2 frames; warps 0 and 1 mix short and long delays -> general loop, warp 2 is all long -> branch-free
    loop), ragged chunk sizes, write-pointer wrap inside a tile, an instance count that is not a multiple of the warp."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    I, T = 84, 50000  # > 48 000: the shared write pointer wraps
    x = orc.noise(12, 0, I, 0, T, 0.5)
    times = [0.0, 1.0, 0.0001, 0.0002, 0.0003, 0.0004, 0.0005, 0.001, 0.0053, 0.0054, 0.0055, 0.0106, 0.0107, 0.0213, 0.0214]  # 0 .. 1 027 samples, and the maximum
    evs = [[(0, loader.EV_HW_CC, 35, times[i] if i < len(times) else 0.05 + 0.9 * i / I), (0, loader.EV_MIDI_CC, 38, (i * 5) % 128)]
           for i in range(I)]
    y = run_bank(kind, x, evs, extra_cuts=(1, 8, 17, 2000, 2001, 4097, 47990))
    name = NAMES[kind]
    for i in list(range(16)) + [31, 32, 63, 64, 70, 83]:
        want = getattr(orc, name)(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, evs[i])
        assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), (env, name, i)


@pytest.mark.parametrize("kind", [fx.DELAY, fx.RACK])
@pytest.mark.parametrize("instances", [6000, 9604])
def test_large_bank_dispatch_is_bit_exact(orc, kind, instances):
    """Banks past one pipeline per SM run 2 (here 188 pipelines) or 4 (301 pipelines, the last CTA three-quarters empty and
    its last pipeline ragged) pipelines per CTA: sampled instances against the oracle, the block cut once mid-tile."""
    T = 1500
    x = orc.noise(21, 0, instances, 0, T, 0.5)
    b = FxBank(kind, instances)
    times = np.linspace(0.02, 1.0, instances, dtype=np.float32)
    b.UpdateHardwareControl(35, times, first=0, count=instances)
    b.UpdateMidiControl(38, 90)
    dx = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    dy = torch.empty_like(dx)
    b.ProcessDevice(dx.data_ptr(), dy.data_ptr(), 1001)
    b.ProcessDevice(dx[1001:].data_ptr(), dy[1001:].data_ptr(), T - 1001)
    torch.cuda.synchronize()
    assert b.LastKernel() == "k_fx_split"
    b.close()
    y = dy.cpu().numpy()
    for i in [0, 31, 32, 63, 64, 95, 96, 127, 128, 2999, 5999, instances - 5, instances - 1]:
        ev = [(0, loader.EV_HW_CC, 35, float(times[i])), (0, loader.EV_MIDI_CC, 38, 90)]
        want = getattr(orc, NAMES[kind])(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, ev)
        assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), (NAMES[kind], instances, i)


def test_default_kernel_is_the_tma_pipeline():
    b = FxBank(fx.RACK, 256)
    x = torch.zeros((64, 256), dtype=torch.float32, device="cuda")
    b.ProcessDevice(x.data_ptr(), x.data_ptr(), 64)
    torch.cuda.synchronize()
    assert b.LastKernel() == "k_fx_split"  # the ring kinds: one warp per pipeline stage
    b.close()
    b = FxBank(fx.FILTER, 256)
    b.ProcessDevice(x.data_ptr(), x.data_ptr(), 64)
    torch.cuda.synchronize()
    assert b.LastKernel() == "k_fx_tma"
    b.close()
    b = FxBank(fx.RACK, 255)  # not a multiple of 4: no tensor map, falls to the register-pipelined kernel
    x = torch.zeros((64, 255), dtype=torch.float32, device="cuda")
    b.ProcessDevice(x.data_ptr(), x.data_ptr(), 64)
    torch.cuda.synchronize()
    assert b.LastKernel() == "k_fx_pf"
    b.close()


def test_plate_rack_bank_with_reverb_controls(orc):
    """OLB_FX_RACK_PLATE: FxRack<1> with the Dattorro plate in the reverb slot, the wiring
    sketched in modules/fxlib/ReverbFx.cpp:11-37.  Per-instance reverb controls through the rack's hardware-control map
    (Fx.h:353-391), ragged chunks, and a re-Init; checked against the composition DelayFx -> plate (left) in ReverbFx's
    dry/wet -> FilterFx -> master volume of pieces that are each pinned on their own."""
    NAN = float("nan")
    I, T = 36, 2600
    x = orc.noise(14, 0, I, 0, T, 0.5)
    b = FxBank(fx.RACK_PLATE, I)
    dec = (0.2 + 0.6 * np.arange(I) / I).astype(np.float32)
    bal = (0.05 + 0.9 * np.arange(I) / I).astype(np.float32)
    cut = (0.1 + 0.8 * ((np.arange(I) * 7) % I) / I).astype(np.float32)
    b.UpdateHardwareControl(32, dec)   # CC_REVERB_TIME -> decay_time -> DattorroVerb_setDecay
    b.UpdateHardwareControl(34, bal)   # CC_REVERB_BALANCE
    b.UpdateHardwareControl(33, cut)   # CC_REVERB_CUTOFF -> cutoff (Hz) -> damping = cutoff / 20 000
    b.UpdateHardwareControl(53, 0.7)   # CC_REVERB_PREFILTER
    b.UpdateHardwareControl(7, 0.6)    # CC_CTL_VOLUME
    dx = torch.from_numpy(x).cuda()
    dy = torch.empty_like(dx)

    def run():
        a = 0
        for n in (1, 999, 1600):
            b.ProcessDevice(dx[a:].data_ptr(), dy[a:].data_ptr(), n)
            a += n
        torch.cuda.synchronize()
        return dy.cpu().numpy()

    y = run()
    b.Init(48000.0)  # clears delay line, filters and plate; keeps every control
    y2 = run()
    assert np.array_equal(bits(y), bits(y2))
    for i in (0, 1, 17, 35):
        d = orc.delayfx(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, [])[:, 0]
        damping = orc.scale(orc.scale(float(cut[i]), 0.0, 1.0, 0.0, 20000.0, 1.0), 0.0, 20000.0, 0.0, 1.0, 1.0)
        wl, _ = orc.dattorro(d, [NAN, 0.7, 0.5, 0.5, float(dec[i]), 0.5, damping])
        mix = wl * bal[i] + d * (np.float32(1.0) - bal[i])
        want = orc.filterfx(np.ascontiguousarray(mix.reshape(-1, 1)), 48000.0, [])[:, 0] * np.float32(0.6)
        assert np.array_equal(bits(y[:, i]), bits(want)), i
    b.close()


# ---- stereo instances: the reference's <2> templates (olb_fx2_*) ----------------------------------------------------------
def run_bank2(kind, x, events_per_instance, sr=48000.0, extra_cuts=(), prefill=0.0, host=False, in_place=False, mode=None):
    """x: [frames][instances][2].  Same control-script convention as run_bank."""
    T, I, _ = x.shape
    b = FxBank(kind, I, sample_rate=sr, channels=2)
    if mode is not None:
        b.SetMode(mode)
    cuts = sorted(c for c in ({0, T} | {e[0] for ev in events_per_instance for e in ev} | set(extra_cuts)) if c <= T)  # later events never fire
    dx = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    dy = dx if in_place else torch.full_like(dx, prefill)
    hy = np.full_like(x, prefill)
    kernels = set()
    for a, z in zip(cuts[:-1], cuts[1:]):
        for i, ev in enumerate(events_per_instance):
            for (f, k, cc, v) in ev:
                if f != a:
                    continue
                if k == loader.EV_MIDI_CC:
                    b.UpdateMidiControl(cc, int(v), first=i, count=1)
                elif k == loader.EV_HW_CC:
                    b.UpdateHardwareControl(cc, v, first=i, count=1)
                elif k == loader.EV_INIT:
                    b.Init(v)
        if host:
            b.ProcessHost(x[a:z], out=hy[a:z])
        else:
            b.ProcessDevice(dx[a:].data_ptr(), dy[a:].data_ptr(), z - a)
        kernels.add(b.LastKernel())
    torch.cuda.synchronize()
    b.close()
    return (hy if host else dy.cpu().numpy()), kernels


@pytest.mark.parametrize("kind", [fx.FILTER, fx.DELAY, fx.REVERB, fx.RACK])
@pytest.mark.parametrize("host", [False, True])
def test_stereo_unit_matches_reference_golden(orc, kind, host):
    """tests/golden/fx.npz stereo entries: the reference's own <2> class over 4 000 stereo frames with the FX_EVENTS script;
    bystander instances with other settings against the oracle."""
    g = np.load(os.path.join(G, "fx.npz"))
    seed, frames, scale = g["stereo_meta"]
    frames = int(frames)
    x2 = orc.noise(int(seed), 0, 2, 0, frames, float(scale))
    other = orc.noise(6, 2, 6, 0, frames, 0.4)
    x = np.concatenate([x2, other], axis=1).reshape(frames, 4, 2)
    evs = [mg.FX_EVENTS, mg.FX2_SHORT_EVENTS, [(0, loader.EV_HW_CC, 35, 0.0013), (100, loader.EV_MIDI_CC, 43, 60)], [(0, loader.EV_HW_CC, 35, 0.9999)]]
    prefill = 0.25 if kind == fx.FILTER else 0.0
    y, _ = run_bank2(kind, x, evs, extra_cuts=(1, 777), prefill=prefill, host=host)
    name = NAMES[kind]
    want = g["fxrack2_full" if kind == fx.RACK else f"{name}2"]
    assert np.array_equal(bits(y[:, 0, :]), bits(want))
    if kind in (fx.DELAY, fx.RACK):  # the reference's own short-delay golden (the wet path and the feedback are heard)
        xs = np.ascontiguousarray(x[:, [0, 0, 2, 3], :])
        ys, _ = run_bank2(kind, xs, [evs[1], evs[1], evs[2], evs[3]], prefill=prefill, host=host)
        assert np.array_equal(bits(ys[:, 0, :]), bits(g[f"{name}2_short"]))
    for i in (1, 2, 3):
        kw = {"prefill": prefill} if kind == fx.FILTER else {}
        w = getattr(orc, name)(np.ascontiguousarray(x[:, i, :]), 48000.0, evs[i], **kw)
        assert np.array_equal(bits(y[:, i, :]), bits(w)), (name, i)


def test_stereo_delay_both_channel1_kernels_and_in_place(orc):
    """DelayFx<2>'s channel 1: a bank whose delays are all long takes the time-parallel kernel, one short delay sends the
    bank to the kernel that walks time; both bit-equal to the oracle, also in place and across a mid-stream re-Init."""
    I, T = 36, 9000
    x = orc.noise(12, 0, 2 * I, 0, T, 0.5).reshape(T, I, 2)
    for short in (False, True):
        evs = []
        for i in range(I):
            t = 0.2 + 0.6 * i / I
            if short and i == 7:
                t = 0.0004
            evs.append([(0, loader.EV_HW_CC, 35, t), (0, loader.EV_MIDI_CC, 36, 40 + i), (0, loader.EV_MIDI_CC, 39, 30 + 2 * i),
                        (5000, loader.EV_INIT, 0, 44100.0)])
        y, kernels = run_bank2(fx.DELAY, x.copy(), evs, in_place=True)
        assert kernels == ({"k_fx2_delay1"} if short else {"k_fx2_put"}), kernels
        for i in (0, 7, 20, 35):
            w = orc.delayfx(np.ascontiguousarray(x[:, i, :]), 48000.0, evs[i])
            assert np.array_equal(bits(y[:, i, :]), bits(w)), (short, i)


def test_stereo_rack_wide_bank_channel0_is_the_mono_rack(orc):
    """2 052 stereo racks (the mono bank inside takes its TMA warp pipeline): channel 0 equals the mono bank on channel 0,
    channel 1 is zero, sampled instances equal the oracle's FxRack<2>."""
    I, T = 2052, 3000
    x = orc.noise(13, 0, 2 * I, 0, T, 0.5).reshape(T, I, 2)
    times = (0.05 + 0.9 * np.arange(I) / I).astype(np.float32)
    b2 = FxBank(fx.RACK, I, channels=2)
    b1 = FxBank(fx.RACK, I)
    for b in (b1, b2):
        b.UpdateHardwareControl(35, times)
    dx = torch.from_numpy(x).cuda()
    dy = torch.empty_like(dx)
    b2.ProcessDevice(dx.data_ptr(), dy.data_ptr(), T)
    m = dx[:, :, 0].contiguous()
    my = torch.empty_like(m)
    b1.ProcessDevice(m.data_ptr(), my.data_ptr(), T)
    torch.cuda.synchronize()
    assert b1.LastKernel() == "k_fx_split"
    y = dy.cpu().numpy()
    assert np.array_equal(bits(y[:, :, 0]), bits(my.cpu().numpy()))
    assert not y[:, :, 1].any()
    for i in (0, 1000, 2051):
        w = orc.fxrack(np.ascontiguousarray(x[:, i, :]), 48000.0, [(0, loader.EV_HW_CC, 35, float(times[i]))])
        assert np.array_equal(bits(y[:, i, :]), bits(w)), i
    b1.close()
    b2.close()


@pytest.mark.parametrize("kind", [fx.FILTER, fx.DELAY, fx.RACK])
def test_stereo_bank_with_an_odd_instance_count(orc, kind):
    """7 stereo instances: the mono bank inside cannot use its tensor maps (instances % 4 != 0) and takes the plain-load kernel;
    short and long delays mixed, so DelayFx<2>'s channel 1 walks time."""
    I, T = 7, 5000
    x = orc.noise(14, 0, 2 * I, 0, T, 0.5).reshape(T, I, 2)
    evs = [[(0, loader.EV_HW_CC, 35, [0.0004, 0.3, 0.9, 0.05, 0.5, 0.0, 1.0][i]), (0, loader.EV_MIDI_CC, 36, 60 + 5 * i), (2000, loader.EV_MIDI_CC, 45, 30 + 10 * i)]
           for i in range(I)]
    prefill = 0.25 if kind == fx.FILTER else 0.0
    y, _ = run_bank2(kind, x, evs, prefill=prefill)
    for i in range(I):
        kw = {"prefill": prefill} if kind == fx.FILTER else {}
        w = getattr(orc, NAMES[kind])(np.ascontiguousarray(x[:, i, :]), 48000.0, evs[i], **kw)
        assert np.array_equal(bits(y[:, i, :]), bits(w)), (kind, i)


@pytest.mark.parametrize("kind", [fx.FILTER, fx.DELAY, fx.REVERB, fx.RACK])
def test_cc_maps_sweep_every_controller_number(orc, kind):
    """Row a10 on the library's side: libolbank restates the CC -> parameter maps in csrc/host_params.h.  Instance i of a
    128-instance bank receives controller number i -- mapped or not -- with a spread of MIDI and hardware values; every instance
    against the oracle (which tests/test_oracle.py pins to the reference's own classes over the same sweep)."""
    T = 400
    x1 = orc.noise(61, 0, 1, 0, T, 0.5)
    x = np.ascontiguousarray(np.repeat(x1, 128, axis=1))
    for evk, values in ((loader.EV_MIDI_CC, (0, 1, 17, 64, 100, 127)), (loader.EV_HW_CC, (0.0, 0.013, 0.5, 1.0))):
        for v in values:
            evs = [[(0, loader.EV_MIDI_CC, 35, 3), (100, evk, cc, v)] for cc in range(128)]
            y = run_bank(kind, x, evs)
            for cc in range(128):
                want = getattr(orc, NAMES[kind])(x1, 48000.0, evs[cc])
                assert np.array_equal(bits(y[:, cc]), bits(want[:, 0])), (NAMES[kind], cc, evk, v)


# ---- the fused tier: olb_fx_set_mode(OLB_MODE_FMA) ---------------------------------------------------------------------------
@pytest.fixture
def fused(orc):
    """The oracle switched to the fused forms (orc_fx_set_mode) for one test."""
    orc.fx_mode(True)
    yield orc
    orc.fx_mode(False)


FMA_VARIANTS = [{"OLB_FX_KERNEL": "simple"}, {"OLB_FX_KERNEL": "pf"}, {"OLB_FX_KERNEL": "tma"},
                {"OLB_FX_KERNEL": "tma", "OLB_FX_SPLIT": "0", "OLB_FX_TMA_CFG": "32x4"}, {"OLB_FX_KERNEL": "tma", "OLB_FX_SPLIT": "0", "OLB_FX_TMA_CFG": "64x8"},
                {"OLB_FX_KERNEL": "tma", "OLB_FX_SPLIT_G": "1", "OLB_FX_TMA_CFG": "32x4"}, {"OLB_FX_KERNEL": "tma", "OLB_FX_SPLIT_G": "2", "OLB_FX_TMA_CFG": "64x4"},
                {"OLB_FX_KERNEL": "tma", "OLB_FX_SPLIT_G": "4", "OLB_FX_TMA_CFG": "32x4"}]


@pytest.mark.parametrize("env", FMA_VARIANTS, ids=lambda e: "-".join(e.values()))
@pytest.mark.parametrize("kind", [fx.FILTER, fx.DELAY, fx.REVERB, fx.RACK])
def test_fma_tier_every_kernel_variant_is_bit_exact_against_the_fused_oracle(fused, monkeypatch, env, kind):
    """OLB_MODE_FMA is a DEFINED arithmetic, not "whatever the compiler contracts": every kernel variant computes the bits of
    the oracle's fused forms (oracle/olb_oracle.cpp svf_half / ring_read / delay_tick / reverb_tick with fast = 1).  Same bank
    as the EXACT variant test -- delays either side of the look-ahead thresholds, ragged cuts, pointer wrap -- with resonance,
    drive and every filter type in play so that each fused operation carries signal."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    I, T = 84, 50000
    x = fused.noise(13, 0, I, 0, T, 0.5)
    times = [0.0, 1.0, 0.0001, 0.0002, 0.0003, 0.0004, 0.0005, 0.001, 0.0053, 0.0054, 0.0055, 0.0106, 0.0107, 0.0213, 0.0214]
    evs = [[(0, loader.EV_HW_CC, 35, times[i] if i < len(times) else 0.05 + 0.9 * i / I), (0, loader.EV_MIDI_CC, 38, (i * 5) % 128),
            (0, loader.EV_MIDI_CC, 36, 20 + i), (0, loader.EV_MIDI_CC, 39, 127 - i), (0, loader.EV_MIDI_CC, 40, (i * 3) % 100), (0, loader.EV_MIDI_CC, 41, (i * 7) % 128),
            (0, loader.EV_MIDI_CC, 42, 30 + i), (0, loader.EV_MIDI_CC, 43, 50 + (i % 60)), (0, loader.EV_MIDI_CC, 45, (i * 11) % 128), (0, loader.EV_MIDI_CC, 46, i)]
           for i in range(I)]
    y = run_bank(kind, x, evs, extra_cuts=(1, 8, 17, 2000, 2001, 4097, 47990), mode=MODE_FMA)
    name = NAMES[kind]
    for i in list(range(16)) + [31, 32, 63, 64, 70, 83]:
        want = getattr(fused, name)(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, evs[i])
        assert np.array_equal(bits(y[:, i]), bits(want[:, 0])), (env, name, i)


def test_fma_tier_is_another_arithmetic_and_exact_is_untouched_by_the_switch(orc):
    """Not vacuous, and reversible: on a resonant rack the FMA tier's bits differ from EXACT's; switching back mid-stream
    continues in EXACT arithmetic from the FMA tier's state (the oracle does the same switch at the same frame)."""
    I, T, cut = 40, 6000, 2500
    x = orc.noise(14, 0, I, 0, T, 0.5)
    ev = [(0, loader.EV_HW_CC, 35, 0.004), (0, loader.EV_MIDI_CC, 36, 90), (0, loader.EV_MIDI_CC, 39, 80), (0, loader.EV_MIDI_CC, 40, 90), (0, loader.EV_MIDI_CC, 41, 64)]
    y_exact = run_bank(fx.RACK, x, [ev] * I, mode=MODE_EXACT)
    y_fma = run_bank(fx.RACK, x, [ev] * I, mode=MODE_FMA)
    assert np.array_equal(bits(y_exact[:, 3]), bits(orc.fxrack(np.ascontiguousarray(x[:, 3:4]), 48000.0, ev)[:, 0]))
    assert not np.array_equal(bits(y_exact), bits(y_fma))
    b = FxBank(fx.RACK, I)
    for (_, k, cc, v) in ev:
        b.UpdateHardwareControl(cc, v) if k == loader.EV_HW_CC else b.UpdateMidiControl(cc, int(v))
    dx = torch.from_numpy(x).cuda()
    dy = torch.empty_like(dx)
    b.SetMode(MODE_FMA)
    b.ProcessDevice(dx.data_ptr(), dy.data_ptr(), cut)
    b.SetMode(MODE_EXACT)
    b.ProcessDevice(dx[cut:].data_ptr(), dy[cut:].data_ptr(), T - cut)
    torch.cuda.synchronize()
    y = dy.cpu().numpy()
    assert np.array_equal(bits(y[:cut]), bits(y_fma[:cut]))
    assert not np.array_equal(bits(y[cut:]), bits(y_exact[cut:])) and not np.array_equal(bits(y[cut:]), bits(y_fma[cut:]))
    np.testing.assert_allclose(y[cut:], y_exact[cut:], rtol=0, atol=1e-4)
    L = _lib.lib()
    assert L.olb_fx_set_mode(b.h, 2) == _lib.E_ARG and L.olb_fx_set_mode(b.h, -1) == _lib.E_ARG and L.olb_fx_set_mode(None, 0) == _lib.E_ARG
    b.close()
    p = FxBank(fx.RACK_PLATE, 2)
    assert L.olb_fx_set_mode(p.h, MODE_FMA) == _lib.E_ARG and L.olb_fx_set_mode(p.h, MODE_EXACT) == _lib.OK
    p.close()


E, H = loader.EV_MIDI_CC, loader.EV_HW_CC
FMA_CORNERS = {
    "default": [],
    "script": mg.FX_EVENTS,
    "resonant": [(0, H, 35, 0.05), (0, E, 36, 120), (0, E, 38, 100), (0, E, 39, 110), (0, E, 40, 100), (0, E, 41, 60), (0, E, 42, 90), (0, E, 43, 90)],
    "all controllers at maximum": [(0, H, 35, 0.0005)] + [(0, E, c, 127) for c in (36, 38, 39, 40, 41, 42, 43, 34)],
    "middle": [(0, H, 35, 0.3)] + [(0, E, c, 64) for c in (36, 38, 39, 40, 41, 42, 43)],
}
FMA_TOL = 1e-4  # of max(1, peak of the EXACT output); include/olbank.h states it, DESIGN.md section 6 has the measured figures


@pytest.mark.parametrize("kind", [fx.FILTER, fx.DELAY, fx.REVERB, fx.RACK])
def test_fma_tier_stays_within_its_stated_tolerance_over_ten_seconds(fused, kind):
    """The tolerance olbank.h states for OLB_MODE_FMA, measured on the device: 480 000 frames of 0.9-scale noise through five
    corners of the CC maps (one instance each, 11 bystanders), FMA against EXACT from the same bank type.  The worst corner
    -- every controller at 127: feedback at its maximum, the loop peaks in the thousands -- is where the two roundings
    drift apart most (about 2e-5 of the peak); anywhere the output stays below 16 they agree to 5e-7."""
    T, I = 480000, 16
    names = list(FMA_CORNERS)
    x = np.repeat(fused.noise(91, 0, 1, 0, T, 0.9), I, axis=1)
    evs = [FMA_CORNERS[names[i % len(names)]] for i in range(I)]
    y_exact = run_bank(kind, x, evs, mode=MODE_EXACT)
    y_fma = run_bank(kind, x, evs, mode=MODE_FMA)
    assert np.isfinite(y_exact).all() and np.isfinite(y_fma).all()
    worst = {}
    for i, n in enumerate(names):
        peak = float(np.abs(y_exact[:, i]).max())
        d = float(np.abs(y_fma[:, i] - y_exact[:, i]).max())
        worst[n] = (d / max(1.0, peak), peak)
        assert d <= FMA_TOL * max(1.0, peak), (NAMES[kind], n, d, peak)
        if peak < 16.0:
            assert d <= 2e-6 * max(1.0, peak), (NAMES[kind], n, d, peak)
    print(f"\n{NAMES[kind]} |FMA - EXACT| / max(1, peak) over 10 s:", {k: f"{v[0]:.2e} (peak {v[1]:.1f})" for k, v in worst.items()})
    i = names.index("resonant")  # and the long stream still equals the fused oracle bit for bit
    want = getattr(fused, NAMES[kind])(np.ascontiguousarray(x[:, i:i + 1]), 48000.0, evs[i])
    assert np.array_equal(bits(y_fma[:, i]), bits(want[:, 0]))


@pytest.mark.parametrize("kind", [fx.FILTER, fx.DELAY, fx.REVERB, fx.RACK])
def test_fma_tier_stereo_banks(fused, kind):
    """olb_fx2_set_mode: both channels in the fused forms, through both of DelayFx<2>'s channel-1 kernels, the static-scratch
    entry point included."""
    I, T = 36, 9000
    x = fused.noise(15, 0, 2 * I, 0, T, 0.5).reshape(T, I, 2)
    for short in ((False, True) if kind == fx.DELAY else (False,)):
        evs = []
        for i in range(I):
            t = 0.0004 if (short and i == 7) else 0.2 + 0.6 * i / I
            evs.append([(0, H, 35, t), (0, E, 36, 40 + i), (0, E, 39, 30 + 2 * i), (0, E, 40, 20 + 2 * i), (0, E, 41, 10 * (i % 12)), (0, E, 42, 60), (0, E, 43, 70),
                        (5000, loader.EV_INIT, 0, 44100.0)])
        prefill = 0.25 if kind == fx.FILTER else 0.0
        y, kernels = run_bank2(kind, x.copy(), evs, prefill=prefill, mode=MODE_FMA)
        if kind == fx.DELAY:
            assert kernels == ({"k_fx2_delay1"} if short else {"k_fx2_put"}), kernels
        for i in (0, 7, 20, 35):
            kw = {"prefill": prefill} if kind == fx.FILTER else {}
            w = getattr(fused, NAMES[kind])(np.ascontiguousarray(x[:, i, :]), 48000.0, evs[i], **kw)
            assert np.array_equal(bits(y[:, i, :]), bits(w)), (NAMES[kind], short, i)
    b = FxBank(kind, I, channels=2)
    b.SetMode(MODE_FMA)
    for (_, k, cc, v) in evs[0][:-1]:
        b.UpdateHardwareControl(cc, v) if k == H else b.UpdateMidiControl(cc, int(v))
    dx = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    dy = torch.full_like(dx, 0.25 if kind == fx.FILTER else 0.0)
    b.ProcessDeviceStatic(dx.data_ptr(), dy.data_ptr(), T)
    torch.cuda.synchronize()
    kw = {"prefill": 0.25} if kind == fx.FILTER else {}
    w = getattr(fused, NAMES[kind])(np.ascontiguousarray(x[:, 5, :]), 48000.0, evs[0][:-1], **kw)
    assert np.array_equal(bits(dy[:, 5, :].cpu().numpy()), bits(w))
    b.close()
