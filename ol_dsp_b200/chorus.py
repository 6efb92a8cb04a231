"""ChorusBank: host-side handle of a bank of mono choruses over the C ABI.  The reference's
chorus is the RNBO patch modules/rnbo/patcher/mono-chorus.rnbopat; the parameter names and 0..1 ranges are that patch's
`param` boxes.  All compute happens in libolbank.so on the GPU."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check

RATE, DEPTH, CUTOFF, Q, MIX, PHASE, PITCH = range(7)
PARAMS = {"rate": RATE, "depth": DEPTH, "cutoff": CUTOFF, "q": Q, "mix": MIX, "phase": PHASE, "pitch": PITCH}


class ChorusBank:
    def __init__(self, instances, sample_rate=48000.0, device=0):
        self.L = _lib.lib()
        self.instances, self.device = int(instances), int(device)
        h = C.c_void_p()
        check(self.L.olb_chorus_create(device, instances, sample_rate, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.L.olb_chorus_destroy(self.h)
            self.h = None

    __del__ = close

    def Init(self, sample_rate):
        check(self.L.olb_chorus_init(self.h, sample_rate))

    def Set(self, param, value, first=0, count=None):
        """param: name or index; value: scalar for the whole range or one value per instance (0..1, clipped; pitch: 0..3 Hz,
        the phasor frequency of the gen~ pitch shifter in front of the delay)."""
        param = PARAMS[param] if isinstance(param, str) else int(param)
        count = self.instances - first if count is None else count
        v = np.ascontiguousarray(value, dtype=np.float32).reshape(-1)
        assert v.size in (1, count)
        check(self.L.olb_chorus_set_params(self.h, first, count, param, v.ctypes.data_as(C.POINTER(C.c_float)), 0 if v.size == 1 else 1))

    def ProcessDevice(self, in_ptr, out_ptr, frames, stream=0):
        check(self.L.olb_chorus_process(self.h, C.c_void_p(in_ptr), C.c_void_p(out_ptr), frames, C.c_void_p(stream)))

    def LastKernel(self):
        return self.L.olb_chorus_last_kernel(self.h).decode()

    def ProcessHost(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        assert x.size == x.shape[0] * self.instances
        out = np.empty_like(x)
        check(self.L.olb_chorus_process_host(self.h, C.c_void_p(x.ctypes.data), C.c_void_p(out.ctypes.data), x.shape[0]))
        return out


class StereoChorusBank:
    """modules/rnbo/patcher/stereo-chorus.rnbopat: per stereo instance two mono choruses, the left with cycle~ phase 1, the
    right with phase 0, every other param shared.  Audio [frames][instances][2] is the memory of a mono bank of 2 x instances
    units (unit 2i left, 2i + 1 right), which is what this class sets up."""

    def __init__(self, instances, sample_rate=48000.0, device=0):
        self.instances = int(instances)
        self.bank = ChorusBank(2 * self.instances, sample_rate, device)
        self.bank.Set(PHASE, np.tile(np.asarray([1.0, 0.0], dtype=np.float32), self.instances))

    def close(self):
        self.bank.close()

    def Init(self, sample_rate):
        self.bank.Init(sample_rate)

    def Set(self, param, value, first=0, count=None):
        """One value per stereo instance (or a scalar): both channels get it.  `phase` is fixed by the patch."""
        assert param not in ("phase", PHASE), "stereo-chorus.rnbopat fixes the phases (left 1, right 0)"
        count = self.instances - first if count is None else count
        v = np.ascontiguousarray(value, dtype=np.float32).reshape(-1)
        self.bank.Set(param, v if v.size == 1 else np.repeat(v, 2), first=2 * first, count=2 * count)

    def ProcessDevice(self, in_ptr, out_ptr, frames, stream=0):
        self.bank.ProcessDevice(in_ptr, out_ptr, frames, stream)

    def ProcessHost(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        assert x.shape[1:] == (self.instances, 2)
        return self.bank.ProcessHost(x.reshape(x.shape[0], 2 * self.instances)).reshape(x.shape)

    def LastKernel(self):
        return self.bank.LastKernel()
