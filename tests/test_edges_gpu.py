"""Edge cases of every bank through the C ABI: empty blocks, one unit, one frame at a time, rejected arguments.  The
reference's own tests are of this kind (test/fx_test.cpp:25-55 processes ONE frame after Init; test/synth_test.cpp checks a
default-constructed voice); a block API adds the empty block and the range checks."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import loader
from ol_dsp_b200 import (MODE_EXACT, BiquadBank, ChorusBank, DattorroBank, FxBank, StereoChorusBank, VoiceBank, _lib, c2_coefs, fx)

pytestmark = pytest.mark.gpu
NAN = float("nan")


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()


def test_empty_blocks_are_no_ops_and_leave_the_state_alone(orc):
    """frames == 0 returns OK without touching buffers (NULL allowed) or state: a stream with empty blocks in it equals the
    stream without them."""
    T = 900
    x = orc.noise(71, 0, 4, 0, T, 0.5)
    k = c2_coefs(4)
    b = BiquadBank(4, 8)
    b.SetMode(MODE_EXACT)
    b.SetCoeffs(k)
    dx, dy = dev(x), torch.empty((T, 4), dtype=torch.float32, device="cuda")
    b.ProcessDevice(0, 0, 0)
    b.ProcessDevice(dx.data_ptr(), dy.data_ptr(), 400)
    b.ProcessDevice(dx[400:].data_ptr(), dy[400:].data_ptr(), 0)
    b.ProcessDevice(dx[400:].data_ptr(), dy[400:].data_ptr(), T - 400)
    torch.cuda.synchronize()
    assert np.array_equal(bits(dy.cpu().numpy()), bits(orc.biquad(x, k, mode=loader.MODE_EXACT)))
    b.close()
    for kind, name in ((fx.DELAY, "delayfx"), (fx.RACK, "fxrack")):
        for ch in (1, 2):
            f = FxBank(kind, 4, channels=ch)
            xs = x if ch == 1 else np.ascontiguousarray(orc.noise(72, 0, 8, 0, T, 0.5).reshape(T, 4, 2))
            dxs, dys = dev(xs), torch.zeros(xs.shape, dtype=torch.float32, device="cuda")
            f.ProcessDevice(0, 0, 0)
            f.ProcessDevice(dxs.data_ptr(), dys.data_ptr(), 500)
            f.ProcessDevice(0, 0, 0)
            f.ProcessDevice(dxs[500:].data_ptr(), dys[500:].data_ptr(), T - 500)
            torch.cuda.synchronize()
            got = dys.cpu().numpy()
            want = getattr(orc, name)(np.ascontiguousarray(xs[:, 2:3]) if ch == 1 else np.ascontiguousarray(xs[:, 2, :]), 48000.0, [])
            assert np.array_equal(bits(got[:, 2] if ch == 1 else got[:, 2, :]), bits(want[:, 0] if ch == 1 else want)), (name, ch)
            f.close()
    c = ChorusBank(4)
    dxc, dyc = dev(x), torch.empty((T, 4), dtype=torch.float32, device="cuda")
    c.ProcessDevice(0, 0, 0)
    c.ProcessDevice(dxc.data_ptr(), dyc.data_ptr(), T)
    c.ProcessDevice(0, 0, 0)
    torch.cuda.synchronize()
    assert np.array_equal(bits(dyc[:, 1].cpu().numpy()), bits(orc.chorus(x[:, 1])))
    c.close()
    d = DattorroBank(4)
    dl, dr = torch.empty((T, 4), dtype=torch.float32, device="cuda"), torch.empty((T, 4), dtype=torch.float32, device="cuda")
    d.ProcessDevice(0, 0, 0, 0)
    d.ProcessDevice(dxc.data_ptr(), dl.data_ptr(), dr.data_ptr(), T)
    torch.cuda.synchronize()
    wl, wr = orc.dattorro(x[:, 3], None)
    assert np.array_equal(bits(dl[:, 3].cpu().numpy()), bits(wl)) and np.array_equal(bits(dr[:, 3].cpu().numpy()), bits(wr))
    d.close()
    v = VoiceBank(4)
    v.ProcessDevice(0)  # no frames, no buffers
    v.close()


def test_one_unit_one_frame_at_a_time(orc):
    """The reference's own call shape -- one object, Process(frame_in, frame_out) per frame -- through the device entry points:
    a bank of ONE unit fed single-frame blocks equals the oracle's block."""
    T = 130
    x = orc.noise(73, 0, 1, 0, T, 0.5)
    dx = dev(x)
    for kind, name in ((fx.FILTER, "filterfx"), (fx.DELAY, "delayfx"), (fx.REVERB, "reverbfx"), (fx.RACK, "fxrack")):
        f = FxBank(kind, 1)
        f.UpdateHardwareControl(35, 0.0005)  # a 24-sample delay, heard within the block
        dy = torch.empty((T, 1), dtype=torch.float32, device="cuda")
        for n in range(T):
            f.ProcessDevice(dx[n:].data_ptr(), dy[n:].data_ptr(), 1)
        torch.cuda.synchronize()
        want = getattr(orc, name)(x, 48000.0, [(0, loader.EV_HW_CC, 35, 0.0005)])
        assert np.array_equal(bits(dy.cpu().numpy()), bits(want)), name
        f.close()
    c = ChorusBank(1)
    c.Set("pitch", 2.0)
    dy = torch.empty((T, 1), dtype=torch.float32, device="cuda")
    for n in range(T):
        c.ProcessDevice(dx[n:].data_ptr(), dy[n:].data_ptr(), 1)
    torch.cuda.synchronize()
    assert np.array_equal(bits(dy[:, 0].cpu().numpy()), bits(orc.chorus(x[:, 0], pitch=2.0)))
    c.close()
    s = StereoChorusBank(1)
    x2 = orc.noise(74, 0, 2, 0, T, 0.5).reshape(T, 1, 2)
    y2 = s.ProcessHost(x2)
    assert np.array_equal(bits(y2[:, 0, 0]), bits(orc.chorus(x2[:, 0, 0], phase=1.0))) and np.array_equal(bits(y2[:, 0, 1]), bits(orc.chorus(x2[:, 0, 1], phase=0.0)))
    s.close()
    d = DattorroBank(1)
    dl, dr = torch.empty((T, 1), dtype=torch.float32, device="cuda"), torch.empty((T, 1), dtype=torch.float32, device="cuda")
    for n in range(T):
        d.ProcessDevice(dx[n:].data_ptr(), dl[n:].data_ptr(), dr[n:].data_ptr(), 1)
    torch.cuda.synchronize()
    wl, wr = orc.dattorro(x[:, 0], None)
    assert np.array_equal(bits(dl[:, 0].cpu().numpy()), bits(wl)) and np.array_equal(bits(dr[:, 0].cpu().numpy()), bits(wr))
    d.close()
    k = c2_coefs(1)
    b = BiquadBank(1, 8)
    b.SetMode(MODE_EXACT)
    b.SetCoeffs(k)
    dy = torch.empty((T, 1), dtype=torch.float32, device="cuda")
    for n in range(T):
        b.ProcessDevice(dx[n:].data_ptr(), dy[n:].data_ptr(), 1)
    torch.cuda.synchronize()
    assert np.array_equal(bits(dy.cpu().numpy()), bits(orc.biquad(x, k, mode=loader.MODE_EXACT)))
    b.close()


def test_bad_arguments_are_rejected_with_a_message():
    L = _lib.lib()
    h = C.c_void_p()
    for create, args in ((L.olb_fx_create, (0, 9, 4, 48000.0)), (L.olb_fx_create, (0, fx.RACK, 0, 48000.0)), (L.olb_fx2_create, (0, fx.RACK_PLATE, 4, 48000.0)),
                         (L.olb_chorus_create, (0, 0, 48000.0)), (L.olb_chorus_create, (0, 4, -1.0)), (L.olb_dattorro_create, (0, 0)),
                         (L.olb_voice_create, (0, 0, 48000.0))):
        assert create(*args, C.byref(h)) == _lib.E_ARG and L.olb_last_error(), create
    f = FxBank(fx.RACK, 4)
    one = np.zeros(1, np.float32)
    assert L.olb_fx_update_midi(f.h, 3, 2, 35, 64) == _lib.E_ARG                      # range [3, 5) of 4 instances
    assert L.olb_fx_update_hw(f.h, 0, 4, 35, None, 0) == _lib.E_ARG                   # NULL values
    assert L.olb_fx_process(f.h, None, None, 16, None) == _lib.E_ARG                  # NULL buffers with frames > 0
    assert L.olb_fx_process(f.h, C.c_void_p(256), C.c_void_p(256), 1 << 32, None) == _lib.E_ARG
    f.close()
    c = ChorusBank(4)
    assert L.olb_chorus_set_params(c.h, 0, 4, 7, one.ctypes.data_as(C.POINTER(C.c_float)), 0) == _lib.E_ARG  # unknown parameter
    assert L.olb_chorus_set_params(c.h, 2, 3, 0, one.ctypes.data_as(C.POINTER(C.c_float)), 0) == _lib.E_ARG  # range
    c.Set("pitch", 99.0)  # clipped to the patch's @max 3, like every RNBO param
    c.close()
    v = VoiceBank(4)
    ev = (_lib.lib().olb_voice_process_events)
    from ol_dsp_b200.voice import VoiceEvent
    arr = (VoiceEvent * 2)()
    arr[0].frame, arr[0].kind, arr[0].a, arr[0].v, arr[0].voice = 10, 2, 60, 100.0, 9   # voice 9 of 4
    assert ev(v.h, C.cast(arr, C.c_void_p), 1, None, None, 1, 0, 1.0, 64, None) == _lib.E_ARG
    arr[0].voice, arr[0].kind = 0, 77                                                   # unknown kind
    assert ev(v.h, C.cast(arr, C.c_void_p), 1, None, None, 1, 0, 1.0, 64, None) == _lib.E_ARG
    v.close()
