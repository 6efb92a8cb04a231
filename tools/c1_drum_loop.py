"""BASELINE.json configs[0] (C1): workouts/drum_loop.wav through the reference's own fx chain ON THE CPU, in the dev
container (the 3.2 MB file is part of /root/reference and does not travel to the GPU box; C1 is a CPU configuration).

The chain is the reference's FxRack<1> (delay -> reverb -> filter -> master volume, modules/fxlib/Fx.h:397-493) with its
default members, Init(44100), compiled from the reference's own headers (oracle/_ref/libolref.so: oracle/ref_drivers.cpp
+ the DaisySP restatement shim, -O2 -ffp-contract=off), fed channel 0 of the file; FxRack<2> is fed both channels.
Prints one JSON object: frames/s, x real time, sha256 of the float32 output -- and the same input through the CPU oracle
restatement (bit-equal, so the digest pins both).

    python tools/c1_drum_loop.py [/root/reference/workouts/drum_loop.wav]
"""
import hashlib
import json
import os
import sys
import time
import wave

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import loader  # noqa: E402


def read_wav(path):
    with wave.open(path, "rb") as w:
        ch, width, sr, n = w.getnchannels(), w.getsampwidth(), w.getframerate(), w.getnframes()
        raw = w.readframes(n)
    assert width == 3, "drum_loop.wav is 24-bit PCM"
    b = np.frombuffer(raw, dtype=np.uint8).reshape(-1, 3).astype(np.int32)
    v = b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16)
    v = np.where(v & 0x800000, v - (1 << 24), v)
    return (v.astype(np.float32) / np.float32(8388608.0)).reshape(n, ch), sr   # miniaudio's s24 -> f32 scaling (1 / 2^23)


def timed(fn, reps=5):
    best, out = 1e30, None
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        best = min(best, time.perf_counter() - t0)
    return best, out


if __name__ == "__main__":
    path = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/workouts/drum_loop.wav"
    x, sr = read_wav(path)
    ref, orc = loader.reference(), loader.oracle()
    assert ref is not None, "oracle/_ref/libolref.so is not built (needs /root/reference)"
    mono = np.ascontiguousarray(x[:, :1])
    res = {"file": os.path.basename(path), "format": f"{x.shape[1]} ch, 24-bit, {sr} Hz, {x.shape[0]} frames = {x.shape[0] / sr:.2f} s",
           "cpu": "1 thread of this container", "build": "g++ -O2 -ffp-contract=off, reference headers compiled unmodified (oracle/Makefile)"}
    for name, inp in (("FxRack<1> on channel 0", mono), ("FxRack<2> on both channels", np.ascontiguousarray(x))):
        dt, y = timed(lambda: ref.fxrack(inp, float(sr)))
        dto, yo = timed(lambda: orc.fxrack(inp, float(sr)))
        res[name] = {"seconds": dt, "frames_per_s": x.shape[0] / dt, "x_real_time": x.shape[0] / sr / dt,
                     "sha256_f32_output": hashlib.sha256(np.ascontiguousarray(y, dtype=np.float32).tobytes()).hexdigest(),
                     "oracle_restatement_bit_equal": bool(np.array_equal(y.view(np.uint32), yo.view(np.uint32))),
                     "oracle_x_real_time": x.shape[0] / sr / dto, "peak": float(np.abs(y).max()), "finite": bool(np.isfinite(y).all())}
    dt, y = timed(lambda: ref.delayfx(mono, float(sr)))
    res["DelayFx<1> on channel 0 (test/fx_test.cpp:25-55's unit)"] = {
        "seconds": dt, "frames_per_s": x.shape[0] / dt, "x_real_time": x.shape[0] / sr / dt,
        "sha256_f32_output": hashlib.sha256(np.ascontiguousarray(y, dtype=np.float32).tobytes()).hexdigest(), "finite": bool(np.isfinite(y).all())}
    dt, (l, r) = timed(lambda: ref.dattorro(np.ascontiguousarray(mono[:, 0])))
    res["DattorroVerb (libs/dattorro-verb/verb.cpp) on channel 0"] = {
        "seconds": dt, "frames_per_s": x.shape[0] / dt, "x_real_time": x.shape[0] / sr / dt,
        "sha256_f32_output_left": hashlib.sha256(np.ascontiguousarray(l, dtype=np.float32).tobytes()).hexdigest()}
    print(json.dumps(res, indent=1))
