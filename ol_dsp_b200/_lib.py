"""ctypes binding of ol_dsp_b200/libolbank.so (the C ABI declared in include/olbank.h).

There is no fallback of any kind: if the shared library is missing this module raises, and every
create call fails when no sm_100 device is present.
"""
import ctypes as C
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libolbank.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "olbank.h")

OK, E_ARG, E_CUDA, E_NCCL, E_STATE = 0, -1, -2, -3, -4
LAYOUT_FRAMES, LAYOUT_PLANAR = 0, 1
KIND_BIQUAD, KIND_PORT, KIND_LPF = 0, 1, 2
MODE_EXACT, MODE_FMA, MODE_FAST = 0, 1, 2


class OlbError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"libolbank error {code}: {text}")
        self.code = code


_F = C.POINTER(C.c_float)
_lib = None


def declared_symbols():
    """Every function name include/olbank.h declares (used by the ABI test)."""
    src = open(HEADER_PATH).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(olb_[a-z0-9_]+)\s*\(", src)))


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OlbError(E_STATE, f"{LIB_PATH} is not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                                "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, u32, u64, i32, f32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_float
    sig = {
        "olb_last_error": ([], C.c_char_p),
        "olb_version": ([], C.c_char_p),
        "olb_device_count": ([C.POINTER(i32)], i32),
        "olb_host_alloc": ([C.POINTER(vp), C.c_size_t], i32),
        "olb_host_free": ([vp], i32),
        "olb_bind_thread_near": ([i32, C.POINTER(i32)], i32),
        "olb_host_alloc_near": ([i32, C.POINTER(vp), C.c_size_t, C.POINTER(i32)], i32),
        "olb_launch_count": ([], u64),
        "olb_noise_fill": ([vp, u32, u32, u32, u64, u64, f32, i32, vp], i32),
        "olb_biquad_create": ([i32, u32, u32, f32, C.POINTER(vp)], i32),
        "olb_biquad_destroy": ([vp], i32),
        "olb_biquad_init": ([vp, f32], i32),
        "olb_biquad_set_kinds": ([vp, C.POINTER(C.c_uint8), u32], i32),
        "olb_biquad_set_mode": ([vp, i32], i32),
        "olb_biquad_set_coeffs": ([vp, u32, u32, _F], i32),
        "olb_biquad_design": ([i32, C.c_double, C.c_double, C.c_double, C.c_double, _F], i32),
        "olb_biquad_process": ([vp, vp, vp, u64, i32, vp], i32),
        "olb_biquad_process_host": ([vp, vp, vp, u64, i32], i32),
        "olb_biquad_get_state": ([vp, _F], i32),
        "olb_biquad_set_state": ([vp, _F], i32),
        "olb_biquad_last_kernel": ([vp], C.c_char_p),
        "olb_biquad_set_variant": ([vp, i32], i32),
        "olb_fx_create": ([i32, i32, u32, f32, C.POINTER(vp)], i32),
        "olb_fx_destroy": ([vp], i32),
        "olb_fx_init": ([vp, f32], i32),
        "olb_fx_set_mode": ([vp, i32], i32),
        "olb_fx_update_midi": ([vp, u32, u32, C.c_uint8, C.c_uint8], i32),
        "olb_fx_update_hw": ([vp, u32, u32, C.c_uint8, _F, u32], i32),
        "olb_fx_process": ([vp, vp, vp, u64, vp], i32),
        "olb_fx_last_kernel": ([vp], C.c_char_p),
        "olb_fx_process_host": ([vp, vp, vp, u64], i32),
        "olb_fx2_create": ([i32, i32, u32, f32, C.POINTER(vp)], i32),
        "olb_fx2_destroy": ([vp], i32),
        "olb_fx2_init": ([vp, f32], i32),
        "olb_fx2_set_mode": ([vp, i32], i32),
        "olb_fx2_update_midi": ([vp, u32, u32, C.c_uint8, C.c_uint8], i32),
        "olb_fx2_update_hw": ([vp, u32, u32, C.c_uint8, _F, u32], i32),
        "olb_fx2_process": ([vp, vp, vp, u64, vp], i32),
        "olb_fx2_last_kernel": ([vp], C.c_char_p),
        "olb_fx2_process_host": ([vp, vp, vp, u64], i32),
        "olb_fx2_process_static": ([vp, vp, vp, u64, vp], i32),
        "olb_voice_process_host": ([vp, vp, vp, u32, i32, f32, u64], i32),
        "olb_voice_process_events": ([vp, vp, u32, vp, vp, u32, i32, f32, u64, vp], i32),
        "olb_dattorro_process_host": ([vp, vp, vp, vp, u64], i32),
        "olb_chorus_process_host": ([vp, vp, vp, u64], i32),
        "olb_voice_create": ([i32, u32, f32, C.POINTER(vp)], i32),
        "olb_voice_destroy": ([vp], i32),
        "olb_voice_init": ([vp, f32], i32),
        "olb_voice_update_midi": ([vp, u32, u32, C.c_uint8, C.c_uint8], i32),
        "olb_voice_update_hw": ([vp, u32, u32, C.c_uint8, _F, u32], i32),
        "olb_voice_update_config": ([vp, u32, u32, _F], i32),
        "olb_voice_note_on": ([vp, u32, u32, C.POINTER(C.c_uint8), u32, C.c_uint8], i32),
        "olb_voice_note_off": ([vp, u32, u32], i32),
        "olb_voice_set_gate": ([vp, u32, u32, i32], i32),
        "olb_voice_set_frequency": ([vp, u32, u32, _F, u32], i32),
        "olb_voice_playing": ([vp, u32, u32, C.POINTER(C.c_uint8)], i32),
        "olb_voice_process": ([vp, vp, vp, u32, i32, f32, vp, u64, vp], i32),
        "olb_comm_unique_id": ([vp], i32),
        "olb_comm_init": ([i32, i32, i32, vp, C.POINTER(vp)], i32),
        "olb_comm_destroy": ([vp], i32),
        "olb_comm_info": ([vp, C.POINTER(i32), C.POINTER(i32)], i32),
        "olb_voice_bus_allreduce": ([vp, vp, u64, u32, vp], i32),
        "olb_voice_bus_allreduce_async": ([vp, vp, u64, u32, vp, i32], i32),
        "olb_comm_wait": ([vp, i32, vp], i32),
        "olb_comm_sync": ([vp], i32),
        "olb_voice_process_host_comm": ([vp, vp, vp, u32, i32, f32, u64], i32),
        "olb_dattorro_create": ([i32, u32, C.POINTER(vp)], i32),
        "olb_dattorro_destroy": ([vp], i32),
        "olb_dattorro_init": ([vp], i32),
        "olb_dattorro_set_param": ([vp, u32, u32, i32, f32], i32),
        "olb_dattorro_set_params": ([vp, u32, u32, i32, _F, u32], i32),
        "olb_dattorro_process": ([vp, vp, vp, vp, u64, vp], i32),
        "olb_dattorro_last_kernel": ([vp], C.c_char_p),
        "olb_chorus_create": ([i32, u32, f32, C.POINTER(vp)], i32),
        "olb_chorus_destroy": ([vp], i32),
        "olb_chorus_init": ([vp, f32], i32),
        "olb_chorus_set_param": ([vp, u32, u32, i32, f32], i32),
        "olb_chorus_set_params": ([vp, u32, u32, i32, _F, u32], i32),
        "olb_chorus_process": ([vp, vp, vp, u64, vp], i32),
        "olb_chorus_last_kernel": ([vp], C.c_char_p),
        "olb_sampler_create": ([i32, u32, u32, C.POINTER(vp)], i32),
        "olb_sampler_destroy": ([vp], i32),
        "olb_sampler_load_sample": ([vp, _F, u64, C.POINTER(u32)], i32),
        "olb_sampler_set_sample": ([vp, u32, u32, C.POINTER(u32), i32], i32),
        "olb_sampler_init": ([vp, f32], i32),
        "olb_sampler_seek": ([vp, u32, u32, u64], i32),
        "olb_sampler_set_loop_start": ([vp, u32, u32, u64], i32),
        "olb_sampler_set_loop_end": ([vp, u32, u32, u64], i32),
        "olb_sampler_set_play_mode": ([vp, u32, u32, i32], i32),
        "olb_sampler_play": ([vp, u32, u32], i32),
        "olb_sampler_pause": ([vp, u32, u32], i32),
        "olb_sampler_toggle_play": ([vp, u32, u32], i32),
        "olb_sampler_gate": ([vp, u32, u32, i32], i32),
        "olb_sampler_position": ([vp, u32, u32, C.POINTER(u64), C.POINTER(C.c_uint8)], i32),
        "olb_sampler_process": ([vp, vp, u64, vp], i32),
        "olb_sampler_process_host": ([vp, vp, u64], i32),
    }
    for name, (args, res) in sig.items():
        f = getattr(L, name)
        f.argtypes, f.restype = args, res
    _lib = L
    return L


def check(rc):
    if rc != OK:
        raise OlbError(rc, lib().olb_last_error().decode())
    return rc


def device_count():
    n = C.c_int(0)
    check(lib().olb_device_count(C.byref(n)))
    return n.value


def launch_count():
    return int(lib().olb_launch_count())
