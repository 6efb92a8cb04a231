"""DattorroBank: host-side mirror of libs/dattorro-verb/verb.h:5-26 for a bank of plates, over the C ABI.
Method names follow the reference's C functions (DattorroVerb_setDecay -> SetDecay ...).  All compute happens in
libolbank.so on the GPU."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check

PREDELAY, PREFILTER, INPUT_DIFFUSION1, INPUT_DIFFUSION2, DECAY, DECAY_DIFFUSION, DAMPING = range(7)


class DattorroBank:
    def __init__(self, instances, device=0):
        self.L = _lib.lib()
        self.instances, self.device = int(instances), int(device)
        h = C.c_void_p()
        check(self.L.olb_dattorro_create(device, instances, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.L.olb_dattorro_destroy(self.h)
            self.h = None

    __del__ = close

    def Init(self):
        check(self.L.olb_dattorro_init(self.h))

    def Set(self, param, value, first=0, count=None):
        """value: scalar for the whole range, or one value per instance."""
        count = self.instances - first if count is None else count
        if np.isscalar(value):
            check(self.L.olb_dattorro_set_param(self.h, first, count, param, float(value)))
        else:
            v = np.ascontiguousarray(value, dtype=np.float32)
            assert v.shape == (count,)
            check(self.L.olb_dattorro_set_params(self.h, first, count, param, v.ctypes.data_as(C.POINTER(C.c_float)), 1))

    def SetPreDelay(self, v, **kw): self.Set(PREDELAY, v, **kw)
    def SetPreFilter(self, v, **kw): self.Set(PREFILTER, v, **kw)
    def SetInputDiffusion1(self, v, **kw): self.Set(INPUT_DIFFUSION1, v, **kw)
    def SetInputDiffusion2(self, v, **kw): self.Set(INPUT_DIFFUSION2, v, **kw)
    def SetDecay(self, v, **kw): self.Set(DECAY, v, **kw)
    def SetDecayDiffusion(self, v, **kw): self.Set(DECAY_DIFFUSION, v, **kw)
    def SetDamping(self, v, **kw): self.Set(DAMPING, v, **kw)

    def ProcessDevice(self, in_ptr, out_l_ptr, out_r_ptr, frames, stream=0):
        check(self.L.olb_dattorro_process(self.h, C.c_void_p(in_ptr), C.c_void_p(out_l_ptr), C.c_void_p(out_r_ptr), frames,
                                          C.c_void_p(stream)))

    def LastKernel(self):
        return self.L.olb_dattorro_last_kernel(self.h).decode()
