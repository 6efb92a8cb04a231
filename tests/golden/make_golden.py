"""Generates tests/golden/*.npz from the COMPILED REFERENCE (oracle/_ref/libolref.so, built from
/root/reference by oracle/Makefile).  Run in the dev container only:

    python tests/golden/make_golden.py

The fixtures are small: short runs are stored whole, long runs as sha256 digests plus sparse samples.
Inputs are the replayable Philox noise (seed in each file), so tests regenerate them.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import loader  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
NAN = float("nan")

DATTORRO_CASES = {
    # name: (params7 or None, seed, frames, scale)
    "default_short": (None, 11, 4096, 0.25),
    "default_wrap": (None, 12, 70000, 0.25),  # crosses the uint16 sample counter wrap and 34 excursion steps
    "long_decay": ([0.3, 0.6, 0.7, 0.5, 0.9, 0.6, 0.8], 13, 20000, 0.25),
    "short_predelay": ([0.0, NAN, NAN, NAN, 0.2, NAN, 0.5], 14, 6000, 0.5),
}

FX_EVENTS = [
    (0, loader.EV_MIDI_CC, 35, 20),     # CC_DELAY_TIME -> 7559 samples
    (0, loader.EV_MIDI_CC, 36, 90),     # CC_DELAY_FEEDBACK
    (3000, loader.EV_MIDI_CC, 45, 40),  # CC_FX_FILTER_CUTOFF
    (3000, loader.EV_HW_CC, 46, 0.6),   # CC_FX_FILTER_RESONANCE
    (9000, loader.EV_MIDI_CC, 39, 80),  # CC_DELAY_BALANCE
    (12000, loader.EV_MIDI_CC, 7, 100),  # CC_VOLUME
]

# stereo runs are 4 000 frames: a delay short enough for the wet path (and the feedback) to be heard several times over
FX2_SHORT_EVENTS = [
    (0, loader.EV_HW_CC, 35, 0.0113),   # CC_DELAY_TIME -> 542.4 samples (fractional read)
    (0, loader.EV_MIDI_CC, 36, 100),    # CC_DELAY_FEEDBACK
    (0, loader.EV_MIDI_CC, 39, 90),     # CC_DELAY_BALANCE
    (1500, loader.EV_MIDI_CC, 37, 30),  # CC_DELAY_CUTOFF
    (2500, loader.EV_HW_CC, 35, 0.004),  # a shorter delay mid-stream
]

VOICE_EVENTS = [
    (0, loader.EV_MIDI_CC, 41, 70), (0, loader.EV_MIDI_CC, 42, 30), (0, loader.EV_MIDI_CC, 73, 64),
    (0, loader.EV_MIDI_CC, 75, 60), (0, loader.EV_MIDI_CC, 108, 10), (0, loader.EV_MIDI_CC, 111, 40),
    (10, loader.EV_NOTE_ON, 60, 100), (2400, loader.EV_NOTE_OFF, 60, 0), (3000, loader.EV_NOTE_ON, 72, 90),
    (3000, loader.EV_MIDI_CC, 5, 30), (5200, loader.EV_NOTE_OFF, 72, 0),
]


# VoiceMap scenario (modules/synthlib/VoiceMap.h): a pad-style kit.  Channel 3 is set twice, so its CCs reach voice 4 only;
# voice 5 is never mapped; note 61 and note 200 are not mapped and must be ignored.
VOICEMAP_MAPPING = [(0, 36, 0), (1, 38, 1), (2, 42, 2), (3, 60, 3), (3, 62, 4)]
VOICEMAP_EVENTS = [(0, loader.EV_MIDI_CC, 41, 40, 1), (0, loader.EV_MIDI_CC, 42, 90, 3), (0, loader.EV_NOTE_ON, 36, 100),
                   (3, loader.EV_NOTE_ON, 38, 90), (400, loader.EV_NOTE_ON, 62, 127), (401, loader.EV_NOTE_ON, 61, 100),
                   (900, loader.EV_NOTE_OFF, 36, 0), (1000, loader.EV_NOTE_ON, 60, 80), (1200, loader.EV_NOTE_ON, 200, 100),
                   (1500, loader.EV_MIDI_CC, 41, 100, 0), (1700, loader.EV_NOTE_ON, 36, 110), (2000, loader.EV_NOTE_OFF, 38, 0),
                   (2400, loader.EV_NOTE_ON, 42, 64), (2600, loader.EV_MIDI_CC, 41, 20, 7)]


# Sample playback scenarios (modules/synthlib/Sample.cpp, SampleSoundSource.h): seed -> a random sample and a random script
# of gates / seeks / loop points (some past the end of the sample) / mode changes / play toggles.
def sampler_scenario(seed, channels, sample_frames=200, frames=1500, n_events=40):
    L = loader
    r = np.random.default_rng(seed)
    data = r.standard_normal((sample_frames, channels)).astype(np.float32)
    kinds = [L.EV_GATE_ON, L.EV_GATE_OFF, L.EV_SAMPLE_SEEK, L.EV_SAMPLE_LOOP_START, L.EV_SAMPLE_LOOP_END, L.EV_SAMPLE_MODE,
             L.EV_SAMPLE_TOGGLE, L.EV_SAMPLE_PLAY, L.EV_SAMPLE_PAUSE]
    ev = []
    for f in sorted(r.integers(0, frames, n_events)):
        k = kinds[r.integers(0, len(kinds))]
        a = int(r.integers(0, sample_frames + 30)) if k in (L.EV_SAMPLE_SEEK, L.EV_SAMPLE_LOOP_START, L.EV_SAMPLE_LOOP_END) else int(r.integers(0, 2))
        ev.append((int(f), k, a, 0.0))
    return data, ev, frames


SAMPLER_CASES = [(seed, 1 + seed % 2) for seed in range(24)]
SAMPLER_FILL = 7.0  # what the output holds before the player runs: frames it does not play must keep it


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    ref = loader.reference()
    orc = loader.oracle()
    assert ref is not None, "oracle/_ref/libolref.so missing: run `make -C oracle` with /root/reference present"

    out = {}
    for name, (params, seed, frames, scale) in DATTORRO_CASES.items():
        x = orc.noise(seed, 0, 1, 0, frames, scale)[:, 0]
        l, r = ref.dattorro(x, params)
        out[f"{name}_meta"] = np.array([seed, frames, scale], dtype=np.float64)
        out[f"{name}_params"] = np.array(params if params is not None else [NAN] * 7, dtype=np.float32)
        if frames <= 8192:
            out[f"{name}_l"], out[f"{name}_r"] = l, r
        else:
            out[f"{name}_sha"] = np.array([sha(l), sha(r)])
            out[f"{name}_tail_l"], out[f"{name}_tail_r"] = l[-256:], r[-256:]
    np.savez_compressed(os.path.join(HERE, "dattorro.npz"), **out)

    fx = {}
    x = orc.noise(21, 0, 1, 0, 16000, 0.5)
    fx["meta"] = np.array([21, 16000, 0.5], dtype=np.float64)
    for name in ("filterfx", "delayfx", "reverbfx", "fxrack"):
        y = getattr(ref, name)(x, 48000.0, FX_EVENTS)
        fx[f"{name}_sha"] = np.array([sha(y)])
        fx[f"{name}_head"] = y[:512, 0]
        fx[f"{name}_tail"] = y[-512:, 0]
    x2 = orc.noise(22, 0, 2, 0, 4000, 0.5)
    fx["stereo_meta"] = np.array([22, 4000, 0.5], dtype=np.float64)
    fx["fxrack2"] = ref.fxrack(x2, 44100.0, FX_EVENTS[:2])
    # the other <2> templates (channel 1: FilterFx<2> leaves frame_out[1] alone -> the 0.25 prefill survives; DelayFx<2>
    # filters channel 0 only; ReverbFx<2> both): Fx.h:88-108, 194-208, 294-300
    fx["filterfx2"] = ref.filterfx(x2, 48000.0, FX_EVENTS, prefill=0.25)
    fx["delayfx2"] = ref.delayfx(x2, 48000.0, FX_EVENTS)
    fx["reverbfx2"] = ref.reverbfx(x2, 48000.0, FX_EVENTS)
    fx["fxrack2_full"] = ref.fxrack(x2, 48000.0, FX_EVENTS)
    fx["delayfx2_short"] = ref.delayfx(x2, 48000.0, FX2_SHORT_EVENTS)
    fx["fxrack2_short"] = ref.fxrack(x2, 48000.0, FX2_SHORT_EVENTS)
    rc, d = ref.test_fx_delay()
    fx["test_fx_delay_nans"] = np.array([rc])
    fx["test_fx_delay_sha"] = np.array([sha(d)])
    fx["test_fx_delay_tail"] = d[-256:]
    np.savez_compressed(os.path.join(HERE, "fx.npz"), **fx)

    v = {}
    y = ref.voice(6000, 48000.0, None, VOICE_EVENTS)
    v["voice"] = y
    p = ref.polyvoice(4, 3000, 48000.0, None,
                      [(0, loader.EV_NOTE_ON, 48, 100), (5, loader.EV_NOTE_ON, 55, 100), (700, loader.EV_NOTE_ON, 64, 100),
                       (1500, loader.EV_NOTE_OFF, 55, 0), (1600, loader.EV_NOTE_ON, 67, 100), (2500, loader.EV_NOTE_OFF, 48, 0)])
    v["poly4"] = p
    v["voicemap"] = ref.voicemap(6, 3000, VOICEMAP_MAPPING, 48000.0, None, VOICEMAP_EVENTS)
    rc, fr = ref.test_synth_voice()
    v["test_synth_voice_failed_mask"] = np.array([rc])
    v["test_synth_voice_frames"] = fr
    sc = [(64, 0, 127, 0, 20000, 1), (24, 0, 127, 0, 1, 1), (0.37, 0, 1, 0, 20000, 1.02), (90, 0, 127, 0, 20000, 2.5),
          (30, 0, 127, 0, 1, 4), (-3, 0, 127, -1, 1, 3), (5, 5, 5, 0, 1, 1)]
    v["scale_args"] = np.array(sc, dtype=np.float32)
    v["scale_out"] = np.array([ref.scale(*a) for a in sc], dtype=np.float32)
    np.savez_compressed(os.path.join(HERE, "synth.npz"), **v)
    sp = {}
    for seed, ch in SAMPLER_CASES:
        data, ev, frames = sampler_scenario(seed, ch)
        o = np.full((frames, ch), SAMPLER_FILL, dtype=np.float32)
        ref.sampler(data, ev, o)
        sp[f"out_{seed}"] = o
    np.savez_compressed(os.path.join(HERE, "sampler.npz"), **sp)
    print("wrote", [f for f in os.listdir(HERE) if f.endswith(".npz")])


if __name__ == "__main__":
    main()
