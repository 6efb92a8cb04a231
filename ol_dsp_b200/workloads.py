"""Synthetic workloads of BASELINE.json: coefficient sets and device noise."""
import ctypes as C

import numpy as np

from . import _lib
from .biquad import design


def c2_coefs(channels, sample_rate=48000.0):
    """8 sections per channel: 4 low-pass (fc log-spaced 200 Hz..12 kHz by channel mod 64, Q .707),
    2 peaking (+3 / -3 dB, Q 2), 1 high-pass 30 Hz, 1 one-pole low-pass as a biquad.  [C][8][5]."""
    base = np.empty((64, 8, 5), dtype=np.float32)
    for m in range(64):
        fc = 200.0 * (12000.0 / 200.0) ** (m / 63.0)
        for s in range(4):
            base[m, s] = design(0, sample_rate, fc, 0.707)
        base[m, 4] = design(2, sample_rate, min(fc * 1.5, 0.45 * sample_rate), 2.0, 3.0)
        base[m, 5] = design(2, sample_rate, max(fc * 0.5, 40.0), 2.0, -3.0)
        base[m, 6] = design(1, sample_rate, 30.0, 0.707)
        base[m, 7] = design(3, sample_rate, min(fc * 2.0, 0.45 * sample_rate), 1.0)
    return np.ascontiguousarray(base[np.arange(channels) % 64])


def noise_device(ptr, seed, first_unit, units, first_frame, frames, scale=1.0, layout=0, stream=0):
    _lib.check(_lib.lib().olb_noise_fill(C.c_void_p(ptr), seed, first_unit, units, first_frame, frames, scale, layout,
                                         C.c_void_p(stream)))
