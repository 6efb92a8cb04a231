"""TEST INFRASTRUCTURE ONLY: ctypes loader for the oracle libraries.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  Nothing under ol_dsp_b200/ does.

  oracle()     -> oracle/libolb_oracle.so   our CPU restatement (built from oracle/olb_oracle.cpp)
  reference()  -> oracle/_ref/libolref.so   the reference's own headers + verb.cpp (built in the dev
                                            container from /root/reference; the .so travels to the GPU box)
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

EV_MIDI_CC, EV_HW_CC, EV_NOTE_ON, EV_NOTE_OFF, EV_INIT, EV_SET_FREQ, EV_GATE_ON, EV_GATE_OFF = range(8)
EV_SAMPLE_SEEK, EV_SAMPLE_LOOP_START, EV_SAMPLE_LOOP_END, EV_SAMPLE_MODE, EV_SAMPLE_TOGGLE, EV_SAMPLE_PLAY, EV_SAMPLE_PAUSE = range(8, 15)
KIND_BIQUAD, KIND_PORT, KIND_LPF = 0, 1, 2
MODE_EXACT, MODE_FMA, MODE_FAST = 0, 1, 2
LAYOUT_FRAMES, LAYOUT_PLANAR = 0, 1


class Event(C.Structure):
    _fields_ = [("frame", C.c_uint32), ("kind", C.c_uint32), ("a", C.c_uint32), ("v", C.c_float), ("voice", C.c_int32)]


def events(lst):
    """lst of (frame, kind, a, v[, voice]) -> (ctypes array, n)"""
    arr = (Event * max(1, len(lst)))()
    for i, e in enumerate(sorted(lst, key=lambda e: e[0])):
        arr[i].frame, arr[i].kind, arr[i].a, arr[i].v = int(e[0]), int(e[1]), int(e[2]), float(e[3])
        arr[i].voice = int(e[4]) if len(e) > 4 else -1
    return arr, len(lst)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def build(force=False):
    """(Re)build libolb_oracle.so, and _ref/libolref.so when the reference tree is present."""
    so = os.path.join(HERE, "libolb_oracle.so")
    src = os.path.join(HERE, "olb_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", HERE, "libolb_oracle.so"], stdout=sys.stderr)
    subprocess.check_call(["make", "-s", "-C", HERE, "ref"], stdout=sys.stderr)  # bench.py's stdout is ONE JSON line


_EVP = C.POINTER(Event)
_F = C.POINTER(C.c_float)
_RUN_SIG = [C.c_int, C.c_float, _EVP, C.c_size_t, _F, _F, C.c_size_t]


class _Fx:
    """Shared wrapper shape for the per-unit run functions of either library."""

    def __init__(self, lib, prefix):
        self.lib, self.prefix = lib, prefix
        for name in ("filterfx", "delayfx", "reverbfx", "fxrack"):
            f = getattr(lib, f"{prefix}_{name}_run")
            f.argtypes, f.restype = _RUN_SIG, C.c_int
        sc = getattr(lib, f"{prefix}_scale")
        sc.argtypes, sc.restype = [C.c_float] * 6, C.c_float
        d = getattr(lib, f"{prefix}_dattorro_run")
        d.argtypes, d.restype = [_F, _F, _F, _F, C.c_size_t], C.c_int

    def scale(self, *a):
        return getattr(self.lib, f"{self.prefix}_scale")(*[float(x) for x in a])

    def _run(self, name, x, sr, ev, prefill=0.0):
        x = np.ascontiguousarray(x, dtype=np.float32)
        n = 1 if x.ndim == 1 else x.shape[1]
        out = np.full_like(x, prefill)
        arr, n_ev = events(ev or [])
        rc = getattr(self.lib, f"{self.prefix}_{name}_run")(n, sr, arr, n_ev, _fp(x), _fp(out), x.shape[0])
        assert rc == 0
        return out

    def fx_mode(self, fast):
        """The fx units' arithmetic: False = the reference's separately rounded operations (default), True = the fused tier
        (oracle only: the compiled reference has no such switch)."""
        if self.prefix == "orc":
            self.lib.orc_fx_set_mode(1 if fast else 0)
        else:
            assert not fast

    def filterfx(self, x, sr=48000.0, ev=None, prefill=0.0):
        return self._run("filterfx", x, sr, ev, prefill)

    def delayfx(self, x, sr=48000.0, ev=None):
        return self._run("delayfx", x, sr, ev)

    def reverbfx(self, x, sr=48000.0, ev=None):
        return self._run("reverbfx", x, sr, ev)

    def fxrack(self, x, sr=48000.0, ev=None):
        return self._run("fxrack", x, sr, ev)

    def dattorro(self, x, params=None):
        x = np.ascontiguousarray(x, dtype=np.float32)
        l, r = np.empty_like(x), np.empty_like(x)
        p = None
        if params is not None:
            p = np.ascontiguousarray(params, dtype=np.float32)
            assert p.shape == (7,)
        rc = getattr(self.lib, f"{self.prefix}_dattorro_run")(_fp(p) if p is not None else None, _fp(x), _fp(l), _fp(r), x.shape[0])
        assert rc == 0
        return l, r


class Oracle(_Fx):
    def __init__(self):
        build()
        lib = C.CDLL(os.path.join(HERE, "libolb_oracle.so"))
        super().__init__(lib, "orc")
        lib.orc_noise_fill.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64, C.c_float, _F, C.c_uint64, C.c_uint64]
        lib.orc_noise_fill.restype = None
        lib.orc_biquad_run.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, _F, _F, _F, _F, C.c_uint64, C.c_int]
        lib.orc_biquad_run.restype = C.c_int
        lib.orc_biquad_run_f64.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p, _F, _F, C.POINTER(C.c_double), C.c_uint64]
        lib.orc_biquad_run_f64.restype = C.c_int
        lib.orc_biquad_fma_coefs.argtypes = [C.c_int, _F, _F]
        lib.orc_biquad_rescale_state.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p, _F, C.c_int, C.c_void_p, _F, C.c_int, _F]
        lib.orc_biquad_rescale_state.restype = C.c_int
        lib.orc_rbj.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, _F]
        lib.orc_rbj.restype = None
        lib.orc_voice_run.argtypes = [C.c_float, _F, _EVP, C.c_size_t, _F, _F, C.c_size_t]
        lib.orc_voice_run.restype = C.c_int
        lib.orc_polyvoice_run.argtypes = [C.c_int, C.c_float, _F, _EVP, C.c_size_t, _F, C.c_size_t]
        lib.orc_polyvoice_run.restype = C.c_int
        lib.orc_dattorro_slab_floats.restype = C.c_uint32
        lib.orc_biquad_bench.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, _F, C.c_uint64, C.c_uint32, C.c_int, C.POINTER(C.c_double)]
        lib.orc_biquad_bench.restype = C.c_double
        lib.orc_biquad_bench_steps.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, _F, C.c_uint64, C.c_uint32, C.c_int, C.c_int, C.c_int,
                                               C.POINTER(C.c_double)]
        lib.orc_biquad_bench_steps.restype = C.c_double
        lib.orc_chorus_params.argtypes = [C.c_float] * 7 + [_F]
        lib.orc_chorus_params.restype = None
        lib.orc_chorus_ring_len.argtypes = [C.c_float]
        lib.orc_chorus_ring_len.restype = C.c_uint32
        lib.orc_chorus_run.argtypes = [_F, C.c_uint32, _F, _F, C.POINTER(C.c_uint32), _F, _F, _F, C.c_uint64, C.c_uint64]
        lib.orc_chorus_run.restype = None
        lib.orc_pitch_hist_len.argtypes = [C.c_float]
        lib.orc_pitch_hist_len.restype = C.c_uint32
        lib.orc_pitch_params.argtypes = [C.c_float, C.c_float, _F]
        lib.orc_pitch_params.restype = None
        lib.orc_pitch_run.argtypes = [_F, C.c_uint32, _F, _F, _F, _F, C.c_uint64, C.c_uint64]
        lib.orc_pitch_run.restype = None

    def pitchshift(self, x, sr=48000.0, pitch=0.0):
        """The gen~ stage alone (pitchshift.gendsp) over x[frames], from rest."""
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
        y = np.empty_like(x)
        hl = int(self.lib.orc_pitch_hist_len(sr))
        par, ph, hist = np.zeros(2, np.float32), np.zeros(1, np.float32), np.zeros(hl, np.float32)
        self.lib.orc_pitch_params(sr, pitch, _fp(par))
        self.lib.orc_pitch_run(_fp(par), hl, _fp(ph), _fp(hist), _fp(x), _fp(y), len(x), 1)
        return y

    def chorus(self, x, sr=48000.0, rate=0.2, depth=0.5, cutoff=0.3, q=0.5, mix=0.5, phase=1.0, pitch=0.0, changes=(), pitch_stage=True):
        """(a9) one chorus instance over x[frames] (defaults = the RNBO patch's param defaults): in~ -> gen~ pitchshift ->
        delay~ -> lores~ -> dry / wet.  changes: list of (frame, dict of params) applied before that frame (state and rings
        kept).  pitch_stage=False leaves the gen~ stage out (delay~ fed by the input itself)."""
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
        y = np.empty_like(x)
        xp = np.empty_like(x)
        cur = dict(rate=rate, depth=depth, cutoff=cutoff, q=q, mix=mix, phase=phase, pitch=pitch)
        par, ppar = np.zeros(7, np.float32), np.zeros(2, np.float32)
        n, hl = int(self.lib.orc_chorus_ring_len(sr)), int(self.lib.orc_pitch_hist_len(sr))
        ring, state, w = np.zeros(n, np.float32), np.zeros(3, np.float32), C.c_uint32(0)
        pph, hist = np.zeros(1, np.float32), np.zeros(hl, np.float32)
        cuts = sorted({0, len(x)} | {f for f, _ in changes})
        for a, z in zip(cuts[:-1], cuts[1:]):
            for f, d in changes:
                if f == a:
                    cur.update(d)
            self.lib.orc_chorus_params(sr, cur["rate"], cur["depth"], cur["cutoff"], cur["q"], cur["mix"], cur["phase"], _fp(par))
            self.lib.orc_pitch_params(sr, cur["pitch"], _fp(ppar))
            if pitch_stage:
                self.lib.orc_pitch_run(_fp(ppar), hl, _fp(pph), _fp(hist), _fp(x[a:]), _fp(xp[a:]), z - a, 1)
            self.lib.orc_chorus_run(_fp(par), n, _fp(state), _fp(ring), C.byref(w), _fp(x[a:]), _fp(xp[a:]) if pitch_stage else None,
                                    _fp(y[a:]), z - a, 1)
        return y

    def sampler(self, data, ev, out, state=None):
        """One sample player (oracle restatement of Sample + SampleSoundSource).  data [sample_frames][channels], out
        [frames][channels] is played INTO (frames not played stay as they are).  Returns the state after the block."""
        data = np.ascontiguousarray(data, dtype=np.float32).reshape(len(data), out.shape[1])
        assert out.dtype == np.float32 and out.flags.c_contiguous
        st = np.zeros(5, dtype=np.uint64) if state is None else np.ascontiguousarray(state, dtype=np.uint64)
        arr, n_ev = events(ev or [])
        fn = self.lib.orc_sampler_run
        fn.argtypes = [C.c_int, _F, C.c_uint64, C.POINTER(C.c_uint64), _EVP, C.c_size_t, _F, C.c_size_t]
        fn.restype = None
        fn(data.shape[1], _fp(data), data.shape[0], st.ctypes.data_as(C.POINTER(C.c_uint64)), arr, n_ev, _fp(out), out.shape[0])
        return st

    def noise(self, seed, first_channel, channels, first_frame, frames, scale=1.0, planar=False):
        """Philox white noise in [-1,1): [frames][channels] (or [channels][frames] when planar)."""
        if planar:
            out = np.empty((channels, frames), dtype=np.float32)
            self.lib.orc_noise_fill(seed, first_channel, channels, first_frame, frames, scale, _fp(out), 1, frames)
        else:
            out = np.empty((frames, channels), dtype=np.float32)
            self.lib.orc_noise_fill(seed, first_channel, channels, first_frame, frames, scale, _fp(out), channels, 1)
        return out

    def rbj(self, typ, fs, f0, q, gain_db=0.0):
        out = np.empty(5, dtype=np.float32)
        self.lib.orc_rbj(typ, fs, f0, q, gain_db, _fp(out))
        return out

    def biquad(self, x, coefs, kinds=None, mode=MODE_EXACT, state=None, planar=False):
        """x: [T][C] (or [C][T] when planar); coefs: [C][S][5]; state: [C][S][4] updated in place."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        coefs = np.ascontiguousarray(coefs, dtype=np.float32)
        ch, sec = coefs.shape[0], coefs.shape[1]
        frames = x.shape[1] if planar else x.shape[0]
        assert (x.shape[0] if planar else x.shape[1]) == ch
        out = np.empty_like(x)
        k = None
        if kinds is not None:
            k = np.ascontiguousarray(kinds, dtype=np.uint8)
            assert k.shape == (sec,)
        if state is not None:
            assert state.dtype == np.float32 and state.shape == (ch, sec, 4) and state.flags.c_contiguous
        rc = self.lib.orc_biquad_run(mode, ch, sec, k.ctypes.data if k is not None else None, _fp(coefs),
                                     _fp(state) if state is not None else None, _fp(x), _fp(out), frames,
                                     LAYOUT_PLANAR if planar else LAYOUT_FRAMES)
        assert rc == 0
        return out

    def biquad_rescale(self, state, coefs_old, mode_old, coefs_new, mode_new, kinds_old=None, kinds_new=None):
        """Re-expresses a FAST history ([C][S][4], in place) in the node scaling of a new coefficient table / mode."""
        co = np.ascontiguousarray(coefs_old, dtype=np.float32)
        cn = np.ascontiguousarray(coefs_new, dtype=np.float32)
        ko = np.ascontiguousarray(kinds_old, dtype=np.uint8) if kinds_old is not None else None
        kn = np.ascontiguousarray(kinds_new, dtype=np.uint8) if kinds_new is not None else ko
        assert state.dtype == np.float32 and state.flags.c_contiguous and state.shape == (co.shape[0], co.shape[1], 4)
        rc = self.lib.orc_biquad_rescale_state(co.shape[0], co.shape[1], ko.ctypes.data if ko is not None else None, _fp(co), mode_old,
                                               kn.ctypes.data if kn is not None else None, _fp(cn), mode_new, _fp(state))
        assert rc == 0

    def biquad_f64(self, x, coefs, kinds=None):
        x = np.ascontiguousarray(x, dtype=np.float32)
        coefs = np.ascontiguousarray(coefs, dtype=np.float32)
        out = np.empty(x.shape, dtype=np.float64)
        k = np.ascontiguousarray(kinds, dtype=np.uint8) if kinds is not None else None
        rc = self.lib.orc_biquad_run_f64(coefs.shape[0], coefs.shape[1], k.ctypes.data if k is not None else None,
                                         _fp(coefs), _fp(x), out.ctypes.data_as(C.POINTER(C.c_double)), x.shape[0])
        assert rc == 0
        return out

    def fma_coefs(self, kind, c5):
        a = np.ascontiguousarray(c5, dtype=np.float32)
        out = np.empty(5, dtype=np.float32)
        self.lib.orc_biquad_fma_coefs(kind, _fp(a), _fp(out))
        return out

    def voice(self, frames, sr=48000.0, config=None, ev=None, trace=False):
        out = np.empty(frames, dtype=np.float32)
        tr = np.empty((frames, 4), dtype=np.float32) if trace else None
        cfg = np.ascontiguousarray(config, dtype=np.float32) if config is not None else None
        arr, n_ev = events(ev or [])
        rc = self.lib.orc_voice_run(sr, _fp(cfg) if cfg is not None else None, arr, n_ev, _fp(out),
                                    _fp(tr) if tr is not None else None, frames)
        assert rc == 0
        return (out, tr) if trace else out

    def polyvoice(self, n, frames, sr=48000.0, config=None, ev=None):
        out = np.empty(frames, dtype=np.float32)
        cfg = np.ascontiguousarray(config, dtype=np.float32) if config is not None else None
        arr, n_ev = events(ev or [])
        rc = self.lib.orc_polyvoice_run(n, sr, _fp(cfg) if cfg is not None else None, arr, n_ev, _fp(out), frames)
        assert rc == 0
        return out

    def biquad_bench(self, coefs, kinds, frames, seed, threads, mode=MODE_EXACT, steps=1, warmup=0):
        """-> (seconds of the `steps` timed passes together, checksum of the last pass's output)"""
        coefs = np.ascontiguousarray(coefs, dtype=np.float32)
        k = np.ascontiguousarray(kinds, dtype=np.uint8) if kinds is not None else None
        chk = C.c_double(0)
        dt = self.lib.orc_biquad_bench_steps(mode, coefs.shape[0], coefs.shape[1], k.ctypes.data if k is not None else None,
                                             _fp(coefs), frames, seed, threads, warmup, steps, C.byref(chk))
        return dt, chk.value


class Reference(_Fx):
    def __init__(self, path):
        lib = C.CDLL(path)
        super().__init__(lib, "ref")
        lib.ref_voice_run.argtypes = [C.c_float, _F, _EVP, C.c_size_t, _F, C.c_size_t]
        lib.ref_voice_run.restype = C.c_int
        lib.ref_polyvoice_run.argtypes = [C.c_int, C.c_float, _F, _EVP, C.c_size_t, _F, C.c_size_t]
        lib.ref_polyvoice_run.restype = C.c_int
        lib.ref_test_fx_delay.argtypes = [_F]
        lib.ref_test_fx_delay.restype = C.c_int
        lib.ref_test_synth_voice.argtypes = [_F]
        lib.ref_test_synth_voice.restype = C.c_int
        lib.ref_describe.restype = C.c_char_p

    def voice(self, frames, sr=48000.0, config=None, ev=None):
        out = np.empty(frames, dtype=np.float32)
        cfg = np.ascontiguousarray(config, dtype=np.float32) if config is not None else None
        arr, n_ev = events(ev or [])
        rc = self.lib.ref_voice_run(sr, _fp(cfg) if cfg is not None else None, arr, n_ev, _fp(out), frames)
        assert rc == 0
        return out

    def polyvoice(self, n, frames, sr=48000.0, config=None, ev=None):
        out = np.empty(frames, dtype=np.float32)
        cfg = np.ascontiguousarray(config, dtype=np.float32) if config is not None else None
        arr, n_ev = events(ev or [])
        rc = self.lib.ref_polyvoice_run(n, sr, _fp(cfg) if cfg is not None else None, arr, n_ev, _fp(out), frames)
        assert rc == 0
        return out

    def voicemap(self, n, frames, mapping, sr=48000.0, config=None, ev=None):
        """The reference's own ol::synth::VoiceMap<1> over n SynthVoices.  mapping = [(channel, note, voice)], fed to
        SetVoice in order; events as everywhere, a MIDI CC's 5th field is the MIDI channel."""
        out = np.empty(frames, dtype=np.float32)
        cfg = np.ascontiguousarray(config, dtype=np.float32) if config is not None else None
        arr, n_ev = events(ev or [])
        m = np.ascontiguousarray(mapping, dtype=np.int32).reshape(-1, 3)
        fn = self.lib.ref_voicemap_run
        fn.argtypes = [C.c_int, C.c_float, _F, C.POINTER(C.c_int32), C.c_size_t, _EVP, C.c_size_t, _F, C.c_size_t]
        fn.restype = C.c_int
        rc = fn(n, sr, _fp(cfg) if cfg is not None else None, m.ctypes.data_as(C.POINTER(C.c_int32)), m.shape[0], arr, n_ev, _fp(out), frames)
        assert rc == 0
        return out

    def sampler(self, data, ev, out):
        """The reference's own ol::synth::Sample + SampleSoundSource over an in-memory data source."""
        data = np.ascontiguousarray(data, dtype=np.float32).reshape(len(data), out.shape[1])
        assert out.dtype == np.float32 and out.flags.c_contiguous
        arr, n_ev = events(ev or [])
        fn = self.lib.ref_sampler_run
        fn.argtypes = [C.c_int, _F, C.c_uint64, _EVP, C.c_size_t, _F, C.c_size_t]
        fn.restype = C.c_int
        assert fn(data.shape[1], _fp(data), data.shape[0], arr, n_ev, _fp(out), out.shape[0]) == 0

    def control_wire(self, which, controller, value):
        """The reference's own ol::ctl::Control + SimpleSerializer::SerializeControl: which 0 MIDI / 1 float / 2 ADC.
        Returns (controller, adc, midi, float bits, wire bytes)."""
        out = (C.c_uint32 * 4)()
        wire = (C.c_uint8 * 64)()
        self.lib.ref_control_wire.argtypes = [C.c_int, C.c_uint32, C.c_double, C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)]
        self.lib.ref_control_wire.restype = C.c_int
        n = self.lib.ref_control_wire(which, controller, float(value), out, wire)
        return out[0], out[1], out[2], out[3], bytes(wire[:n])

    def test_fx_delay(self):
        out = np.empty(48000, dtype=np.float32)
        return self.lib.ref_test_fx_delay(_fp(out)), out

    def test_synth_voice(self):
        fr = np.empty(4, dtype=np.float32)
        return self.lib.ref_test_synth_voice(_fp(fr)), fr


_oracle = None
_reference = None


def oracle():
    global _oracle
    if _oracle is None:
        _oracle = Oracle()
    return _oracle


def reference():
    """The compiled reference, or None when oracle/_ref/libolref.so has not been built."""
    global _reference
    if _reference is None:
        path = os.path.join(HERE, "_ref", "libolref.so")
        if not os.path.exists(path):
            try:
                build()
            except Exception:
                pass
        if os.path.exists(path):
            _reference = Reference(path)
    return _reference
