"""SamplerBank: host-side handle of a bank of sample players over the C ABI -- `instances` x
ol::synth::Sample behind SampleSoundSource (modules/synthlib/Sample.h:16-52, SampleSoundSource.h:14-41) reading from a
device-resident sample pool.  Method names are the reference's; every method takes an instance range.  Process plays INTO
the buffer: frames of instances that are not playing, or are past the end of their sample, stay as they are."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check

ONESHOT, LOOP = 0, 1


class SamplerBank:
    def __init__(self, instances, channels=1, device=0):
        self.L = _lib.lib()
        self.instances, self.channels, self.device = int(instances), int(channels), int(device)
        h = C.c_void_p()
        check(self.L.olb_sampler_create(device, instances, channels, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.L.olb_sampler_destroy(self.h)
            self.h = None

    __del__ = close

    def _range(self, first, count):
        return first, (self.instances - first if count is None else count)

    def LoadSample(self, data):
        """data: [frames] or [frames][channels] floats -> sample id in the pool."""
        d = np.ascontiguousarray(data, dtype=np.float32).reshape(len(data), self.channels)
        sid = C.c_uint32()
        check(self.L.olb_sampler_load_sample(self.h, d.ctypes.data_as(C.POINTER(C.c_float)), d.shape[0], C.byref(sid)))
        return sid.value

    def SetSample(self, sample_id, first=0, count=None):
        """One id for the range, or one per instance.  The instances become fresh Sample objects."""
        first, count = self._range(first, count)
        ids = np.ascontiguousarray(sample_id, dtype=np.uint32).reshape(-1)
        assert ids.size in (1, count)
        check(self.L.olb_sampler_set_sample(self.h, first, count, ids.ctypes.data_as(C.POINTER(C.c_uint32)), 0 if ids.size == 1 else 1))

    def Init(self, sample_rate):
        check(self.L.olb_sampler_init(self.h, sample_rate))

    def Seek(self, frame_index, first=0, count=None):
        check(self.L.olb_sampler_seek(self.h, *self._range(first, count), int(frame_index)))

    def SetLoopStart(self, frame_index, first=0, count=None):
        check(self.L.olb_sampler_set_loop_start(self.h, *self._range(first, count), int(frame_index)))

    def SetLoopEnd(self, frame_index, first=0, count=None):
        check(self.L.olb_sampler_set_loop_end(self.h, *self._range(first, count), int(frame_index)))

    def SetPlayMode(self, mode, first=0, count=None):
        check(self.L.olb_sampler_set_play_mode(self.h, *self._range(first, count), int(mode)))

    def Play(self, first=0, count=None):
        check(self.L.olb_sampler_play(self.h, *self._range(first, count)))

    def Pause(self, first=0, count=None):
        check(self.L.olb_sampler_pause(self.h, *self._range(first, count)))

    def TogglePlay(self, first=0, count=None):
        check(self.L.olb_sampler_toggle_play(self.h, *self._range(first, count)))

    def GateOn(self, first=0, count=None):  # SampleSoundSource::GateOn: Seek(0) + Play()
        check(self.L.olb_sampler_gate(self.h, *self._range(first, count), 1))

    def GateOff(self, first=0, count=None):  # SampleSoundSource::GateOff: Pause()
        check(self.L.olb_sampler_gate(self.h, *self._range(first, count), 0))

    def Position(self, first=0, count=None):
        first, count = self._range(first, count)
        pos, playing = np.zeros(count, dtype=np.uint64), np.zeros(count, dtype=np.uint8)
        check(self.L.olb_sampler_position(self.h, first, count, pos.ctypes.data_as(C.POINTER(C.c_uint64)), playing.ctypes.data_as(C.POINTER(C.c_uint8))))
        return pos, playing

    def ProcessDevice(self, out_ptr, frames, stream=0):
        check(self.L.olb_sampler_process(self.h, C.c_void_p(out_ptr), frames, C.c_void_p(stream)))

    def ProcessHost(self, out):
        """out: float32 [frames][instances * channels], played into in place."""
        assert out.dtype == np.float32 and out.flags.c_contiguous and out.shape[1] == self.instances * self.channels
        check(self.L.olb_sampler_process_host(self.h, C.c_void_p(out.ctypes.data), out.shape[0]))
