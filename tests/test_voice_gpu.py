"""GPU parity of the SynthVoice bank (row a6) and the Polyvoice mixdown (row a8) through the C ABI.
Bit-exact: oscillator output and phase, both envelope traces (stage logic + level recurrence), portamento.
Tolerance (<= 1e-5 of full scale, the bar north_star states): the voice OUTPUT, because Svf::SetFreq's per-sample sinf
runs in CUDA's libm on the device and in glibc in the reference; and the bus, whose summation order differs."""
import os

import numpy as np
import pytest
import torch

from oracle import loader
from ol_dsp_b200 import Polyvoice, VoiceBank, VoiceMap
from tests.golden import make_golden as mg

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-5


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def apply(bank, e, voice=None):
    f, k, a, v = e[:4]
    rng = {} if voice is None else {"first": voice, "count": 1}
    if k == loader.EV_MIDI_CC:
        bank.UpdateMidiControl(a, int(v), **rng)
    elif k == loader.EV_HW_CC:
        bank.UpdateHardwareControl(a, v, **rng)
    elif k == loader.EV_NOTE_ON:
        bank.NoteOn(a, int(v), **rng)
    elif k == loader.EV_NOTE_OFF:
        bank.NoteOff(**rng)
    elif k == loader.EV_INIT:
        bank.Init(v)
    elif k == loader.EV_SET_FREQ:
        bank.SetFrequency(v, **rng)
    elif k == loader.EV_GATE_ON:
        bank.GateOn(**rng)
    elif k == loader.EV_GATE_OFF:
        bank.GateOff(**rng)


def run_voices(V, frames, events_per_voice, config=None, trace=False, sr=48000.0):
    b = VoiceBank(V, sample_rate=sr)
    if config is not None:
        b.UpdateConfig(config)
    cuts = sorted({0, frames} | {e[0] for ev in events_per_voice for e in ev})
    out = torch.empty((frames, V), dtype=torch.float32, device="cuda")
    tr = torch.empty((frames, V, 4), dtype=torch.float32, device="cuda") if trace else None
    for a, z in zip(cuts[:-1], cuts[1:]):
        for i, ev in enumerate(events_per_voice):
            for e in ev:
                if e[0] == a:
                    apply(b, e, i)
        b.ProcessDevice(z - a, out_ptr=out[a:].data_ptr(), trace_ptr=tr[a:].data_ptr() if trace else 0)
    torch.cuda.synchronize()
    b.close()
    return out.cpu().numpy(), (tr.cpu().numpy() if trace else None)


def test_voice_matches_reference_golden_and_traces(orc):
    """tests/golden/synth.npz 'voice': the reference's own SynthVoice run through VOICE_EVENTS (CCs, two notes,
    portamento change).  Voice 1 of the bank runs the script; voices 0 and 2 are bystanders with other scripts."""
    g = np.load(os.path.join(G, "synth.npz"))
    frames = g["voice"].shape[0]
    other = [(0, loader.EV_NOTE_ON, 40, 100), (100, loader.EV_HW_CC, 41, 0.5), (4000, loader.EV_NOTE_OFF, 40, 0)]
    y, tr = run_voices(3, frames, [other, mg.VOICE_EVENTS, []], trace=True)
    assert np.abs(y[:, 1] - g["voice"]).max() <= TOL
    want, wtr = orc.voice(frames, 48000.0, None, mg.VOICE_EVENTS, trace=True)
    assert np.array_equal(bits(tr[:, 1, 0]), bits(wtr[:, 0])), "oscillator output"
    assert np.array_equal(bits(tr[:, 1, 1]), bits(wtr[:, 1])), "amp envelope"
    assert np.array_equal(bits(tr[:, 1, 2]), bits(wtr[:, 2])), "filter envelope"
    assert np.array_equal(bits(tr[:, 1, 3]), bits(wtr[:, 3])), "phase"
    assert np.abs(y[:, 1] - want).max() <= TOL
    w0, t0 = orc.voice(frames, 48000.0, None, other, trace=True)
    assert np.array_equal(bits(tr[:, 0, 3]), bits(t0[:, 3])) and np.abs(y[:, 0] - w0).max() <= TOL
    assert float(np.abs(y[:, 2]).max()) == 0.0  # an idle voice is silent


def test_reference_synth_test_voice_default_constructor():
    """test/synth_test.cpp:102-148 replayed on a one-voice bank: frame == 0 right after NoteOn+NoteOff, != 0 and != 1 on
    the next gated frame, == 0 with amp_env_amount 0 and != 0 with 1."""
    b = VoiceBank(1, sample_rate=128.0)
    f = torch.full((1, 1), 7.0, dtype=torch.float32, device="cuda")

    def frame():
        b.ProcessDevice(1, out_ptr=f.data_ptr())
        torch.cuda.synchronize()
        return float(f[0, 0])

    b.NoteOn(60, 100)
    b.NoteOff()
    assert frame() == 0.0
    b.NoteOn(60, 100)
    v = frame()
    assert v != 0.0 and v != 1.0
    b.UpdateHardwareControl(7, 0.0)  # CC_CTL_VOLUME -> amp_env_amount
    assert frame() == 0.0
    b.UpdateHardwareControl(7, 1.0)
    assert frame() != 0.0
    b.close()


def test_polyvoice_allocation_and_bus(orc):
    """tests/golden/synth.npz 'poly4': Polyvoice with 4 voices, first-free allocation, accumulate into a zeroed bus."""
    g = np.load(os.path.join(G, "synth.npz"))
    want = g["poly4"]
    frames = want.shape[0]
    ev = [(0, loader.EV_NOTE_ON, 48, 100), (5, loader.EV_NOTE_ON, 55, 100), (700, loader.EV_NOTE_ON, 64, 100),
          (1500, loader.EV_NOTE_OFF, 55, 0), (1600, loader.EV_NOTE_ON, 67, 100), (2500, loader.EV_NOTE_OFF, 48, 0)]
    bank = VoiceBank(4)
    poly = Polyvoice(bank)
    bus = torch.zeros((frames, 1), dtype=torch.float32, device="cuda")
    cuts = sorted({0, frames} | {e[0] for e in ev})
    for a, z in zip(cuts[:-1], cuts[1:]):
        for e in ev:
            if e[0] == a:
                (poly.NoteOn if e[1] == loader.EV_NOTE_ON else poly.NoteOff)(e[2], e[3])
        poly.Process(z - a, bus[a:].data_ptr())
    torch.cuda.synchronize()
    assert np.abs(bus.cpu().numpy()[:, 0] - want).max() <= 4 * TOL  # 4 voices summed
    assert list(bank.Playing()) == [0, 67, 64, 0]
    bank.close()


def test_wide_bank_bus_is_deterministic_and_equals_the_sum_of_voices():
    """C4-shaped (scaled down in time): 65 536 voices, note 36 + (v mod 61), presets through the MIDI maps; the stereo bus
    equals the float64 sum of the per-voice outputs within re-association error, and is bit-identical run to run."""
    V, T = 65536, 256
    notes = (36 + np.arange(V) % 61).astype(np.uint8)

    def run():
        b = VoiceBank(V)
        for cc, val in ((41, 70), (42, 30), (73, 64), (75, 60), (108, 10), (111, 40)):
            b.UpdateMidiControl(cc, val)
        b.NoteOn(notes)
        out = torch.empty((T, V), dtype=torch.float32, device="cuda")
        bus = torch.zeros((T, 2), dtype=torch.float32, device="cuda")
        b.ProcessDevice(T, out_ptr=out.data_ptr(), bus_ptr=bus.data_ptr(), bus_channels=2, bus_gain=1.0 / V)
        torch.cuda.synchronize()
        b.close()
        return out, bus

    out, bus = run()
    _, bus2 = run()
    assert torch.equal(bus, bus2)
    assert torch.equal(bus[:, 0], bus[:, 1])
    ref = out.double().sum(dim=1) / V
    assert float((bus[:, 0].double() - ref).abs().max()) <= 1e-6
    assert float(out.abs().max()) > 0.01 and torch.isfinite(out).all()


def test_sample_accurate_event_list(orc):
    """olb_voice_process_events: one call, the event list cuts the block at frame granularity
    and routes notes through Polyvoice's first-free allocation; same golden as the hand-cut test above."""
    from ol_dsp_b200.voice import process_events
    g = np.load(os.path.join(G, "synth.npz"))
    want = g["poly4"]
    frames = want.shape[0]
    ev = [(0, loader.EV_NOTE_ON, 48, 100), (5, loader.EV_NOTE_ON, 55, 100), (700, loader.EV_NOTE_ON, 64, 100),
          (1500, loader.EV_NOTE_OFF, 55, 0), (1600, loader.EV_NOTE_ON, 67, 100), (2500, loader.EV_NOTE_OFF, 48, 0)]
    assert (loader.EV_MIDI_CC, loader.EV_HW_CC, loader.EV_NOTE_ON, loader.EV_NOTE_OFF) == (0, 1, 2, 3)  # same codes as olbank.h
    bank = VoiceBank(4)
    bus = torch.zeros((frames, 1), dtype=torch.float32, device="cuda")
    process_events(bank, ev, frames, bus_ptr=bus.data_ptr(), bus_channels=1, accumulate=True)
    torch.cuda.synchronize()
    assert np.abs(bus.cpu().numpy()[:, 0] - want).max() <= 4 * TOL
    assert list(bank.Playing()) == [0, 67, 64, 0]
    with pytest.raises(Exception):
        process_events(bank, [(10, 2, 60, 100), (5, 2, 62, 100)][::-1] + [(frames + 1, 2, 60, 100)], frames, bus_ptr=bus.data_ptr())
    bank.close()


def test_dense_event_list_on_a_wide_bank(orc):
    """C4's 65 536 voices with a dense sample-accurate script: 64 NoteOn and 32 NoteOff at distinct frames of one 4 800-frame
    block, each addressed to its own voice (a cut uploads the touched voices' parameter columns only, and the retrigger is
    cleared by the voice that consumed it).  The touched voices against the oracle, an untouched one silent."""
    from ol_dsp_b200.voice import process_events
    V, frames = 65536, 4800
    voices = [(i * 1021 + 7) % V for i in range(64)]
    ev = [(13 + 61 * i, loader.EV_NOTE_ON, 40 + i, 100, v) for i, v in enumerate(voices)]
    ev += [(2500 + 53 * i, loader.EV_NOTE_OFF, 40 + 2 * i, 0, voices[2 * i]) for i in range(32)]
    bank = VoiceBank(V)
    out = torch.empty((frames, V), dtype=torch.float32, device="cuda")
    process_events(bank, ev, frames, out_ptr=out.data_ptr())
    torch.cuda.synchronize()
    for i in (0, 1, 2, 31, 62, 63):
        script = [(13 + 61 * i, loader.EV_NOTE_ON, 40 + i, 100)]
        if i % 2 == 0:
            script.append((2500 + 53 * (i // 2), loader.EV_NOTE_OFF, 40 + i, 0))
        want = orc.voice(frames, ev=script)
        got = out[:, voices[i]].cpu().numpy()
        assert np.abs(got - want).max() <= TOL, (i, float(np.abs(got - want).max()))
        assert np.abs(want).max() > 0.01
    assert float(out[:, 5].abs().max()) == 0.0  # a voice nobody played
    playing = bank.Playing()
    assert playing[voices[1]] == 41 and playing[voices[0]] == 0
    bank.close()


def test_event_list_mixed_with_fan_out_controls_and_same_frame_events(orc):
    """Per-voice events ride along with the launch, a CC that fans out to every voice cuts the block, several events of one
    voice may share a frame (NoteOn + CC; NoteOff + NoteOn), an event may sit exactly on a cut or on the block's end, and the
    whole script may be split over two calls: all against the oracle replaying each voice's own script."""
    from ol_dsp_b200.voice import process_events
    V, frames = 96, 6000
    per_voice = {3: [(0, loader.EV_NOTE_ON, 50, 100), (700, loader.EV_MIDI_CC, 41, 90), (700, loader.EV_NOTE_OFF, 50, 0), (700, loader.EV_NOTE_ON, 62, 100)],
                 40: [(1, loader.EV_NOTE_ON, 45, 100), (2000, loader.EV_HW_CC, 42, 0.7), (4999, loader.EV_NOTE_OFF, 45, 0)],
                 95: [(2000, loader.EV_NOTE_ON, 70, 100), (2000, loader.EV_MIDI_CC, 75, 20), (3000, loader.EV_NOTE_OFF, 70, 0), (6000, loader.EV_NOTE_ON, 71, 100)]}
    fan = [(2000, loader.EV_MIDI_CC, 73, 30), (4000, loader.EV_HW_CC, 41, 0.4)]  # voice < 0: every voice
    ev = [(f, k, a, v, voice) for voice, lst in per_voice.items() for (f, k, a, v) in lst] + [(f, k, a, v, -1) for (f, k, a, v) in fan]
    ev.sort(key=lambda t: t[0])  # stable: list order within a frame is kept (per-voice events before the fan-out ones)
    bank = VoiceBank(V)
    out = torch.empty((frames, V), dtype=torch.float32, device="cuda")
    cut = 2500
    first = [t for t in ev if t[0] < cut]
    second = [(t[0] - cut,) + t[1:] for t in ev if t[0] >= cut]
    process_events(bank, first, cut, out_ptr=out.data_ptr())
    process_events(bank, second, frames - cut, out_ptr=out[cut:].data_ptr())
    torch.cuda.synchronize()
    y = out.cpu().numpy()
    for voice, lst in per_voice.items():
        script = sorted([e for e in lst if e[0] < frames] + fan, key=lambda t: t[0])
        want = orc.voice(frames, ev=script)
        assert np.abs(y[:, voice] - want).max() <= TOL, (voice, float(np.abs(y[:, voice] - want).max()))
        assert np.abs(want).max() > 0.01
    assert not y[:, 7].any()
    assert list(bank.Playing()[[3, 40, 95]]) == [62, 0, 71]
    bank.close()


def test_cc_maps_sweep_every_controller_number(orc):
    """SynthVoice's CC maps as restated in csrc/host_params.h: voice i of a 128-voice bank receives controller number i (MIDI and
    hardware values, before and during a note); every voice against the oracle, which tests/test_oracle.py pins to the
    reference's own SynthVoice over the same sweep."""
    from ol_dsp_b200.voice import process_events
    frames = 700
    for kind, values in ((loader.EV_MIDI_CC, (0, 33, 127)), (loader.EV_HW_CC, (0.0, 0.37, 1.0))):
        for v in values:
            ev = [(0, kind, cc, v, cc) for cc in range(128)] + [(5, loader.EV_NOTE_ON, 57, 100, cc) for cc in range(128)]
            ev += [(300, kind, cc, values[1], cc) for cc in range(128)] + [(500, loader.EV_NOTE_OFF, 57, 0, cc) for cc in range(128)]
            bank = VoiceBank(128)
            out = torch.empty((frames, 128), dtype=torch.float32, device="cuda")
            process_events(bank, ev, frames, out_ptr=out.data_ptr())
            torch.cuda.synchronize()
            y = out.cpu().numpy()
            bank.close()
            for cc in range(128):
                want = orc.voice(frames, ev=[(0, kind, cc, v), (5, loader.EV_NOTE_ON, 57, 100), (300, kind, cc, values[1]), (500, loader.EV_NOTE_OFF, 57, 0)])
                assert np.abs(y[:, cc] - want).max() <= TOL, (cc, kind, v, float(np.abs(y[:, cc] - want).max()))


def test_voicemap_against_the_reference_golden():
    """ol::synth::VoiceMap over a bank: tests/golden/synth.npz 'voicemap' is the reference's own
    VoiceMap<1> over six SynthVoices (pads on five notes, channel 3 set twice, unmapped notes / channels ignored)."""
    want = np.load(os.path.join(G, "synth.npz"))["voicemap"]
    frames = want.shape[0]
    bank = VoiceBank(6)
    vm = VoiceMap(bank)
    for ch, note, v in mg.VOICEMAP_MAPPING:
        vm.SetVoice(ch, note, v)
    with pytest.raises(ValueError):
        vm.SetVoice(5, 70, 0)  # voice 0 already sits under note 36
    vm.SetVoice(20, 70, 5)  # channel out of range: ignored like upstream
    assert 70 not in vm.note2voice
    bus = torch.zeros((frames, 1), dtype=torch.float32, device="cuda")
    ev = mg.VOICEMAP_EVENTS
    cuts = sorted({0, frames} | {e[0] for e in ev})
    for a, z in zip(cuts[:-1], cuts[1:]):
        for (f, k, x, v, *ch) in ev:
            if f != a:
                continue
            if k == loader.EV_MIDI_CC:
                vm.UpdateMidiControl(ch[0], x, v)
            elif k == loader.EV_NOTE_ON:
                vm.NoteOn(x, v)
            else:
                vm.NoteOff(x, v)
        vm.Process(z - a, bus[a:].data_ptr())
    torch.cuda.synchronize()
    assert np.abs(bus.cpu().numpy()[:, 0] - want).max() <= 5 * TOL  # five sounding voices summed
    assert list(bank.Playing()) == [36, 0, 42, 60, 62, 0]
    bank.close()
