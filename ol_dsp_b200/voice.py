"""VoiceBank / Polyvoice: host-side mirror of modules/synthlib/SynthVoice.h and Polyvoice.h for a bank of voices, over the
C ABI.  Method names are the reference's.  All DSP runs in libolbank.so on the GPU; the voice-allocation policy of
Polyvoice (first free voice, first voice playing the note: Polyvoice.h:35-51) is host logic and lives here."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check

_F = C.POINTER(C.c_float)
_U8 = C.POINTER(C.c_uint8)


class VoiceBank:
    def __init__(self, voices, sample_rate=48000.0, device=0):
        self.L = _lib.lib()
        self.voices, self.device = int(voices), int(device)
        h = C.c_void_p()
        check(self.L.olb_voice_create(device, voices, sample_rate, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.L.olb_voice_destroy(self.h)
            self.h = None

    __del__ = close

    def _range(self, first, count):
        return first, (self.voices - first if count is None else count)

    def Init(self, sample_rate):
        check(self.L.olb_voice_init(self.h, sample_rate))

    def UpdateMidiControl(self, control, value, first=0, count=None):
        first, count = self._range(first, count)
        check(self.L.olb_voice_update_midi(self.h, first, count, int(control), int(value)))

    def UpdateHardwareControl(self, control, value, first=0, count=None):
        first, count = self._range(first, count)
        v = np.ascontiguousarray(value, dtype=np.float32).reshape(-1)
        assert v.size in (1, count)
        check(self.L.olb_voice_update_hw(self.h, first, count, int(control), v.ctypes.data_as(_F), 0 if v.size == 1 else 1))

    def UpdateConfig(self, config16, first=0, count=None):
        first, count = self._range(first, count)
        c = np.ascontiguousarray(config16, dtype=np.float32)
        assert c.shape == (16,)
        check(self.L.olb_voice_update_config(self.h, first, count, c.ctypes.data_as(_F)))

    def NoteOn(self, note, velocity=100, first=0, count=None):
        """note: one MIDI note for the whole range, or one per voice."""
        first, count = self._range(first, count)
        n = np.ascontiguousarray(note, dtype=np.uint8).reshape(-1)
        assert n.size in (1, count)
        check(self.L.olb_voice_note_on(self.h, first, count, n.ctypes.data_as(_U8), 0 if n.size == 1 else 1, int(velocity)))

    def NoteOff(self, first=0, count=None):
        first, count = self._range(first, count)
        check(self.L.olb_voice_note_off(self.h, first, count))

    def GateOn(self, first=0, count=None):
        first, count = self._range(first, count)
        check(self.L.olb_voice_set_gate(self.h, first, count, 1))

    def GateOff(self, first=0, count=None):
        first, count = self._range(first, count)
        check(self.L.olb_voice_set_gate(self.h, first, count, 0))

    def SetFrequency(self, freq, first=0, count=None):
        first, count = self._range(first, count)
        f = np.ascontiguousarray(freq, dtype=np.float32).reshape(-1)
        assert f.size in (1, count)
        check(self.L.olb_voice_set_frequency(self.h, first, count, f.ctypes.data_as(_F), 0 if f.size == 1 else 1))

    def Playing(self, first=0, count=None):
        first, count = self._range(first, count)
        p = np.zeros(count, dtype=np.uint8)
        check(self.L.olb_voice_playing(self.h, first, count, p.ctypes.data_as(_U8)))
        return p

    def ProcessDevice(self, frames, out_ptr=0, bus_ptr=0, bus_channels=1, accumulate=False, bus_gain=1.0, trace_ptr=0, stream=0):
        check(self.L.olb_voice_process(self.h, C.c_void_p(out_ptr), C.c_void_p(bus_ptr), bus_channels, int(bool(accumulate)),
                                       float(bus_gain), C.c_void_p(trace_ptr), frames, C.c_void_p(stream)))


class VoiceEvent(C.Structure):
    """olb_voice_event (include/olbank.h)."""
    _fields_ = [("frame", C.c_uint32), ("kind", C.c_uint32), ("a", C.c_uint32), ("v", C.c_float), ("voice", C.c_int32)]


EV_MIDI_CC, EV_HW_CC, EV_NOTE_ON, EV_NOTE_OFF = range(4)


def process_events(bank, events, frames, out_ptr=0, bus_ptr=0, bus_channels=1, accumulate=False, bus_gain=1.0, stream=0):
    """Sample-accurate control: events = [(frame, kind, a, v[, voice])], applied before the frame they name (voice < 0 or
    omitted: Polyvoice allocation)."""
    ev = sorted(events, key=lambda e: e[0])
    arr = (VoiceEvent * max(1, len(ev)))()
    for i, e in enumerate(ev):
        arr[i].frame, arr[i].kind, arr[i].a, arr[i].v = int(e[0]), int(e[1]), int(e[2]), float(e[3])
        arr[i].voice = int(e[4]) if len(e) > 4 else -1
    check(bank.L.olb_voice_process_events(bank.h, C.cast(arr, C.c_void_p), len(ev), C.c_void_p(out_ptr), C.c_void_p(bus_ptr), bus_channels,
                                          int(bool(accumulate)), float(bus_gain), frames, C.c_void_p(stream)))


class Polyvoice:
    """ol::synth::Polyvoice (Polyvoice.h:12-86) over a VoiceBank: NoteOn takes the first voice that is not playing,
    NoteOff releases the first voice playing that note, setters fan out to every voice, Process accumulates the voices
    into the caller's (zeroed) bus."""

    def __init__(self, bank):
        self.bank = bank

    def NoteOn(self, note, velocity=100):
        free = np.flatnonzero(self.bank.Playing() == 0)
        if free.size:
            self.bank.NoteOn(note, velocity, first=int(free[0]), count=1)

    def NoteOff(self, note, velocity=0):
        hit = np.flatnonzero(self.bank.Playing() == note)
        if hit.size:
            self.bank.NoteOff(first=int(hit[0]), count=1)

    def UpdateMidiControl(self, control, value):
        self.bank.UpdateMidiControl(control, value)

    def UpdateHardwareControl(self, control, value):
        self.bank.UpdateHardwareControl(control, value)

    def Process(self, frames, bus_ptr, bus_channels=1, stream=0):
        self.bank.ProcessDevice(frames, bus_ptr=bus_ptr, bus_channels=bus_channels, accumulate=True, stream=stream)


class VoiceMap:
    """ol::synth::VoiceMap (modules/synthlib/VoiceMap.h:14-90) over a VoiceBank: a note -> voice table (a pad kit: every
    mapped note triggers ITS voice) and a channel -> voice table for MIDI CCs (the voice last set on that channel).
    Notes and channels that are not mapped are ignored, as in the reference.  Process accumulates the bank into the
    caller's (zeroed) bus; voices that were never triggered are silent.  One deviation: the reference walks its note
    table in Process, so a voice set under two notes would be advanced twice per frame there -- SetVoice refuses that."""

    MAP_SIZE, CHANNELS = 128, 16

    def __init__(self, bank):
        self.bank = bank
        self.note2voice, self.channel2voice = {}, {}

    def SetVoice(self, channel, note, voice):
        if not (0 <= note < self.MAP_SIZE and 0 <= channel < self.CHANNELS):
            return  # VoiceMap.h:46-47: silently ignored
        if not 0 <= voice < self.bank.voices:
            raise ValueError("voice %d is not in the bank" % voice)
        if any(v == voice and n != note for n, v in self.note2voice.items()):
            raise ValueError("voice %d is already mapped under another note" % voice)
        self.note2voice[note] = voice
        self.channel2voice[channel] = voice

    def Init(self, sample_rate):
        self.bank.Init(sample_rate)

    def NoteOn(self, note, velocity=100):
        v = self.note2voice.get(note)
        if v is not None:
            self.bank.NoteOn(note, velocity, first=v, count=1)

    def NoteOff(self, note, velocity=0):
        v = self.note2voice.get(note)
        if v is not None:
            self.bank.NoteOff(first=v, count=1)

    def UpdateMidiControl(self, channel, control, value):
        v = self.channel2voice.get(channel)
        if v is not None:
            self.bank.UpdateMidiControl(control, value, first=v, count=1)

    def Process(self, frames, bus_ptr, bus_channels=1, stream=0):
        self.bank.ProcessDevice(frames, bus_ptr=bus_ptr, bus_channels=bus_channels, accumulate=True, stream=stream)
