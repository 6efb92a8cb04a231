"""BiquadBank: host-side mirror of the reference's object conventions over the C ABI.

Method names follow the reference units (Init / Process / Update-style setters; e.g.
modules/fxlib/Fx.h:83-114).  All compute happens in libolbank.so; this file only marshals.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import LAYOUT_FRAMES, LAYOUT_PLANAR, MODE_EXACT, MODE_FMA, check

_F = C.POINTER(C.c_float)


def _fp(a):
    return a.ctypes.data_as(_F)


def design(typ, fs, f0, q, gain_db=0.0):
    out = np.empty(5, dtype=np.float32)
    check(_lib.lib().olb_biquad_design(typ, fs, f0, q, gain_db, _fp(out)))
    return out


class BiquadBank:
    def __init__(self, channels, sections, sample_rate=48000.0, device=0):
        self.L = _lib.lib()
        self.channels, self.sections, self.device = int(channels), int(sections), int(device)
        h = C.c_void_p()
        check(self.L.olb_biquad_create(device, channels, sections, sample_rate, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.L.olb_biquad_destroy(self.h)
            self.h = None

    __del__ = close

    def Init(self, sample_rate):
        check(self.L.olb_biquad_init(self.h, sample_rate))

    def SetKinds(self, kinds):
        k = np.ascontiguousarray(kinds, dtype=np.uint8)
        check(self.L.olb_biquad_set_kinds(self.h, k.ctypes.data_as(C.POINTER(C.c_uint8)), k.size))

    def SetMode(self, mode):
        check(self.L.olb_biquad_set_mode(self.h, mode))

    def SetVariant(self, variant):
        check(self.L.olb_biquad_set_variant(self.h, variant))

    def SetCoeffs(self, coefs, first=0):
        c = np.ascontiguousarray(coefs, dtype=np.float32)
        assert c.ndim == 3 and c.shape[1:] == (self.sections, 5)
        check(self.L.olb_biquad_set_coeffs(self.h, first, c.shape[0], _fp(c)))

    def ProcessDevice(self, in_ptr, out_ptr, frames, layout=LAYOUT_FRAMES, stream=0):
        """Raw device pointers (ints), e.g. torch tensor.data_ptr()."""
        check(self.L.olb_biquad_process(self.h, C.c_void_p(in_ptr), C.c_void_p(out_ptr), frames, layout, C.c_void_p(stream)))

    def ProcessHostPtr(self, in_ptr, out_ptr, frames, layout=LAYOUT_FRAMES):
        check(self.L.olb_biquad_process_host(self.h, C.c_void_p(in_ptr), C.c_void_p(out_ptr), frames, layout))

    def ProcessBlock(self, x, layout=LAYOUT_FRAMES):
        """x: numpy [frames][channels] (or [channels][frames] when planar) -> same shape."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        frames = x.shape[0] if layout == LAYOUT_FRAMES else x.shape[1]
        assert (x.shape[1] if layout == LAYOUT_FRAMES else x.shape[0]) == self.channels
        out = np.empty_like(x)
        self.ProcessHostPtr(x.ctypes.data, out.ctypes.data, frames, layout)
        return out

    def Process(self, frame_in):
        """One frame for every channel: the reference's per-frame Process(in, out)."""
        return self.ProcessBlock(np.asarray(frame_in, dtype=np.float32).reshape(1, self.channels))[0]

    def GetState(self):
        st = np.empty((self.channels, self.sections, 4), dtype=np.float32)
        check(self.L.olb_biquad_get_state(self.h, _fp(st)))
        return st

    def SetState(self, st):
        st = np.ascontiguousarray(st, dtype=np.float32)
        assert st.shape == (self.channels, self.sections, 4)
        check(self.L.olb_biquad_set_state(self.h, _fp(st)))

    def LastKernel(self):
        return self.L.olb_biquad_last_kernel(self.h).decode()
