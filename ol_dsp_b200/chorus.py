"""ChorusBank: host-side handle of a bank of mono choruses over the C ABI.  The reference's
chorus is the RNBO patch modules/rnbo/patcher/mono-chorus.rnbopat; the parameter names and 0..1 ranges are that patch's
`param` boxes.  All compute happens in libolbank.so on the GPU."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check

RATE, DEPTH, CUTOFF, Q, MIX, PHASE = range(6)
PARAMS = {"rate": RATE, "depth": DEPTH, "cutoff": CUTOFF, "q": Q, "mix": MIX, "phase": PHASE}


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
        """param: name or index; value: scalar for the whole range or one value per instance (0..1, clipped)."""
        param = PARAMS[param] if isinstance(param, str) else int(param)
        count = self.instances - first if count is None else count
        v = np.ascontiguousarray(value, dtype=np.float32).reshape(-1)
        assert v.size in (1, count)
        check(self.L.olb_chorus_set_params(self.h, first, count, param, v.ctypes.data_as(C.POINTER(C.c_float)), 0 if v.size == 1 else 1))

    def ProcessDevice(self, in_ptr, out_ptr, frames, stream=0):
        check(self.L.olb_chorus_process(self.h, C.c_void_p(in_ptr), C.c_void_p(out_ptr), frames, C.c_void_p(stream)))

    def LastKernel(self):
        return self.L.olb_chorus_last_kernel(self.h).decode()
