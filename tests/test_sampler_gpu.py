"""GPU parity of the sample playback bank through the C ABI: bit-exact against
tests/golden/sampler.npz -- outputs of the reference's OWN ol::synth::Sample + SampleSoundSource (Sample.cpp compiled from
where it lies, over an in-memory SampleDataSource) -- and against the oracle restatement on large banks."""
import os

import numpy as np
import pytest
import torch

from oracle import loader
from ol_dsp_b200 import SamplerBank, sampler
from tests.golden import make_golden as mg

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
L = loader


@pytest.fixture(scope="module")
def orc():
    return loader.oracle()


def apply(bank, i, ev):
    (f, k, a, v) = ev[:4]
    one = dict(first=i, count=1)
    if k == L.EV_GATE_ON:
        bank.GateOn(**one)
    elif k == L.EV_GATE_OFF:
        bank.GateOff(**one)
    elif k == L.EV_SAMPLE_SEEK:
        bank.Seek(a, **one)
    elif k == L.EV_SAMPLE_LOOP_START:
        bank.SetLoopStart(a, **one)
    elif k == L.EV_SAMPLE_LOOP_END:
        bank.SetLoopEnd(a, **one)
    elif k == L.EV_SAMPLE_MODE:
        bank.SetPlayMode(sampler.LOOP if a else sampler.ONESHOT, **one)
    elif k == L.EV_SAMPLE_TOGGLE:
        bank.TogglePlay(**one)
    elif k == L.EV_SAMPLE_PLAY:
        bank.Play(**one)
    elif k == L.EV_SAMPLE_PAUSE:
        bank.Pause(**one)


def run_bank(bank, scripts, frames, fill, host=False):
    """scripts[i] = event list of instance i; the block is cut at every event frame."""
    W = bank.instances * bank.channels
    cuts = sorted({0, frames} | {e[0] for ev in scripts for e in ev})
    out = torch.full((frames, W), fill, dtype=torch.float32, device="cuda")
    out_h = np.full((frames, W), fill, dtype=np.float32)
    for a, z in zip(cuts[:-1], cuts[1:]):
        for i, ev in enumerate(scripts):
            for e in ev:
                if e[0] == a:
                    apply(bank, i, e)
        if host:
            bank.ProcessHost(out_h[a:z])
        else:
            bank.ProcessDevice(out[a:].data_ptr(), z - a)
    torch.cuda.synchronize()
    return out_h if host else out.cpu().numpy()


@pytest.mark.parametrize("channels", [1, 2])
@pytest.mark.parametrize("host", [False, True], ids=["device", "host"])
def test_matches_the_reference_golden(channels, host):
    g = np.load(os.path.join(G, "sampler.npz"))
    cases = [(seed, ch) for seed, ch in mg.SAMPLER_CASES if ch == channels]
    bank = SamplerBank(len(cases), channels)
    bank.Init(48000.0)
    scripts, frames = [], None
    for i, (seed, ch) in enumerate(cases):
        data, ev, frames = mg.sampler_scenario(seed, ch)
        bank.SetSample(bank.LoadSample(data), first=i, count=1)
        scripts.append(ev)
    y = run_bank(bank, scripts, frames, mg.SAMPLER_FILL, host=host)
    for i, (seed, ch) in enumerate(cases):
        want = g[f"out_{seed}"]
        assert np.array_equal(y[:, i * ch:(i + 1) * ch].view(np.uint32), want.view(np.uint32)), (seed, ch)
    bank.close()


@pytest.mark.parametrize("channels", [1, 2])
def test_large_bank_against_the_oracle(orc, channels):
    """4 100 instances (not a multiple of the block) over 37 samples of different lengths: one-shots that run out, loops with
    and without loop points, idle instances, long straight runs (the tiled fast path) and the positions afterwards."""
    I, T = 4100, 3000
    r = np.random.default_rng(5)
    bank = SamplerBank(I, channels)
    samples = [r.standard_normal((int(n), channels)).astype(np.float32) for n in r.integers(1, 9000, 37)]
    samples[3] = np.zeros((0, channels), dtype=np.float32)  # an empty sample: Read never returns a frame
    ids = [bank.LoadSample(s) for s in samples]
    which = r.integers(0, len(ids), I)
    bank.SetSample(np.array([ids[w] for w in which], dtype=np.uint32))
    scripts = []
    for i in range(I):
        n = len(samples[which[i]])
        ev = []
        if i % 7 != 0:  # every seventh instance stays idle
            if i % 3 == 0:
                ev += [(0, L.EV_SAMPLE_MODE, 1, 0.0)]
                if i % 2 == 0:
                    a, b = sorted(int(v) for v in r.integers(0, n + 5, 2))
                    ev += [(0, L.EV_SAMPLE_LOOP_START, a, 0.0), (0, L.EV_SAMPLE_LOOP_END, b, 0.0)]
            ev += [(0, L.EV_GATE_ON, 0, 0.0)]
            if i % 5 == 0:
                ev += [(0, L.EV_SAMPLE_SEEK, int(r.integers(0, n + 5)), 0.0)]
        scripts.append(ev)
    # all events sit at frame 0 except two bank-wide cuts, so the loop over instances above stays cheap
    for i in range(0, I, 11):
        scripts[i] = scripts[i] + [(1001, L.EV_SAMPLE_TOGGLE, 0, 0.0), (2017, L.EV_GATE_ON, 0, 0.0)]
    y = run_bank(bank, scripts, T, 3.5)
    pos, playing = bank.Position()
    for i in list(range(0, 300)) + list(range(300, I, 13)) + [I - 1]:
        want = np.full((T, channels), 3.5, dtype=np.float32)
        st = orc.sampler(samples[which[i]], scripts[i], want)
        assert np.array_equal(y[:, i * channels:(i + 1) * channels].view(np.uint32), want.view(np.uint32)), i
        assert int(pos[i]) == int(st[4]) and int(playing[i]) == int(st[0]), i
    bank.close()


def test_argument_errors_are_loud():
    with pytest.raises(Exception):
        SamplerBank(8, channels=3)
    bank = SamplerBank(8)
    with pytest.raises(Exception):
        bank.SetSample(0)  # nothing loaded yet
    sid = bank.LoadSample(np.ones(10, dtype=np.float32))
    with pytest.raises(Exception):
        bank.SetSample(sid, first=4, count=5)
    with pytest.raises(Exception):
        bank.SetPlayMode(2)
    bank.SetSample(sid)
    bank.Play()
    out = torch.zeros((4, 8), dtype=torch.float32, device="cuda")
    bank.ProcessDevice(out.data_ptr(), 4)
    torch.cuda.synchronize()
    assert float(out.sum()) == 32.0
    bank.close()
