"""FxBank: host-side mirror of the reference's fxlib units (modules/fxlib/Fx.h) for a bank of mono instances, over
the C ABI.  Method names are the reference's: Init, UpdateMidiControl, UpdateHardwareControl, Process (block form).
All compute happens in libolbank.so on the GPU; the CC -> parameter maths runs on the host inside the library."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check

FILTER, DELAY, REVERB, RACK, RACK_PLATE = range(5)


class FxBank:
    """channels = 1: a bank of FilterFx<1> / ... (audio [frames][instances]); channels = 2: the reference's <2> templates,
    second-channel behaviour included (audio [frames][instances][2], include/olbank.h olb_fx2_*)."""

    def __init__(self, kind, instances, sample_rate=48000.0, device=0, channels=1):
        self.L = _lib.lib()
        self.kind, self.instances, self.device, self.channels = int(kind), int(instances), int(device), int(channels)
        assert self.channels in (1, 2)
        self._p = "olb_fx_" if self.channels == 1 else "olb_fx2_"
        h = C.c_void_p()
        check(self._f("create")(device, kind, instances, sample_rate, C.byref(h)))
        self.h = h

    def _f(self, name):
        return getattr(self.L, self._p + name)

    def close(self):
        if getattr(self, "h", None):
            self._f("destroy")(self.h)
            self.h = None

    __del__ = close

    def Init(self, sample_rate):
        check(self._f("init")(self.h, sample_rate))

    def SetMode(self, mode):
        """MODE_EXACT (default; the reference's separately rounded arithmetic) or MODE_FMA (the fused tier, olbank.h)."""
        check(self._f("set_mode")(self.h, int(mode)))

    def UpdateMidiControl(self, control, value, first=0, count=None):
        count = self.instances - first if count is None else count
        check(self._f("update_midi")(self.h, first, count, int(control), int(value)))

    def UpdateHardwareControl(self, control, value, first=0, count=None):
        """value: scalar for the whole range, or one value per instance."""
        count = self.instances - first if count is None else count
        v = np.ascontiguousarray(value, dtype=np.float32).reshape(-1)
        assert v.size in (1, count)
        check(self._f("update_hw")(self.h, first, count, int(control), v.ctypes.data_as(C.POINTER(C.c_float)), 0 if v.size == 1 else 1))

    def ProcessDevice(self, in_ptr, out_ptr, frames, stream=0):
        check(self._f("process")(self.h, C.c_void_p(in_ptr), C.c_void_p(out_ptr), frames, C.c_void_p(stream)))

    def ProcessDeviceStatic(self, in_ptr, out_ptr, frames, stream=0):
        """Stereo banks only: ProcessDevice without any device allocation (olb_fx2_process_static: scratch in a static
        device array shared by every call on the device; at most 32 768 instances)."""
        assert self.channels == 2
        check(self.L.olb_fx2_process_static(self.h, C.c_void_p(in_ptr), C.c_void_p(out_ptr), frames, C.c_void_p(stream)))

    def ProcessHost(self, x, out=None):
        """x: host array [frames][instances] (or [frames][instances][2]); returns the processed block (out, if given, is
        filled in place: FilterFx<2> keeps its channel 1)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.empty_like(x) if out is None else out
        assert out.dtype == np.float32 and out.flags.c_contiguous and out.shape == x.shape
        assert x.size == x.shape[0] * self.instances * self.channels
        check(self._f("process_host")(self.h, C.c_void_p(x.ctypes.data), C.c_void_p(out.ctypes.data), x.shape[0]))
        return out

    def LastKernel(self):
        return self._f("last_kernel")(self.h).decode()


def FilterFx(instances, **kw): return FxBank(FILTER, instances, **kw)
def DelayFx(instances, **kw): return FxBank(DELAY, instances, **kw)
def ReverbFx(instances, **kw): return FxBank(REVERB, instances, **kw)
def FxRack(instances, **kw): return FxBank(RACK, instances, **kw)
def FxRackPlate(instances, **kw): return FxBank(RACK_PLATE, instances, **kw)
