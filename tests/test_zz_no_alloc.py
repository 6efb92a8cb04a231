"""olb_fx2_process_static: the stereo fx banks' process call that allocates nothing (scratch in a static device array,
include/olbank.h).  No counterpart in the reference.  Checked through the
C ABI against the existing entry point olb_fx2_process, bit for bit (the tolerance the fx tests use), on two calls in a row
(state -- filters, delay lines, the shared static array -- carries from the first into the second).

This file sorts last on purpose: a failure here must not hide another test."""
import ctypes as C

import numpy as np
import pytest
import torch

from ol_dsp_b200 import FxBank, _lib, fx, noise_device

pytestmark = pytest.mark.gpu


def bits(t):
    return t.cpu().numpy().view(np.uint32)


def _pair(kind, I, times):
    a, b = FxBank(kind, I, channels=2), FxBank(kind, I, channels=2)
    for bank in (a, b):
        bank.UpdateHardwareControl(35, times)  # CC_DELAY_TIME, per instance
        bank.UpdateMidiControl(36, 80)         # CC_DELAY_FEEDBACK
        bank.UpdateMidiControl(45, 50)         # CC_FX_FILTER_CUTOFF
    return a, b


@pytest.mark.parametrize("kind,I,short", [(fx.FILTER, 516, False), (fx.DELAY, 516, False), (fx.DELAY, 516, True), (fx.REVERB, 516, False),
                                          (fx.RACK, 516, False), (fx.RACK, 8192, False)])
def test_static_scratch_entry_matches_process_on_two_calls(kind, I, short):
    calls = [1500, 700] if I <= 1024 else [600, 300]  # 8 192 instances: 256-frame cuts -> several cuts per call
    times = (0.05 + 0.9 * np.arange(I) / I).astype(np.float32)
    if short:
        times[7] = 0.0004  # one short delay sends DelayFx<2>'s channel 1 to the kernel that walks time
    ref, sta = _pair(kind, I, times)
    T = sum(calls)
    x = torch.empty((T, I, 2), dtype=torch.float32, device="cuda")
    noise_device(x.data_ptr(), 31, 0, 2 * I, 0, T, 0.5)
    want = torch.full_like(x, 0.25)  # FilterFx<2> leaves channel 1 of the output as the caller had it
    got = torch.full_like(x, 0.25)
    st = torch.cuda.current_stream().cuda_stream
    t0 = 0
    for n in calls:
        ref.ProcessDevice(x[t0:].data_ptr(), want[t0:].data_ptr(), n, st)
        sta.ProcessDeviceStatic(x[t0:].data_ptr(), got[t0:].data_ptr(), n, st)
        assert sta.LastKernel() == ref.LastKernel()
        torch.cuda.synchronize()
        assert np.array_equal(bits(got[t0:t0 + n]), bits(want[t0:t0 + n])), (kind, I, t0)
        t0 += n
    assert torch.isfinite(got).all() and got[:, :, 0].abs().max() > 0
    ref.close()
    sta.close()


def test_static_scratch_entry_allocates_nothing_and_rejects_oversize():
    I, T = 2048, 512
    b = FxBank(fx.RACK, I, channels=2)
    x = torch.empty((T, I, 2), dtype=torch.float32, device="cuda")
    noise_device(x.data_ptr(), 32, 0, 2 * I, 0, T, 0.5)
    y = torch.empty_like(x)
    b.ProcessDeviceStatic(x.data_ptr(), y.data_ptr(), T)  # first call: symbol address looked up, parameters uploaded
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        b.ProcessDeviceStatic(x.data_ptr(), y.data_ptr(), T)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0  # no device memory taken or returned by the calls
    b.close()
    big = FxBank(fx.FILTER, 32768 + 4, channels=2)
    L = _lib.lib()
    rc = L.olb_fx2_process_static(big.h, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), 1, C.c_void_p(0))
    assert rc == _lib.E_ARG and b"exceed the static scratch" in L.olb_last_error()
    big.close()
