"""Development aid: times biquad-bank kernel shapes in ONE process (CUDA events, resident buffers).

    python tools/sweep_inproc.py [skew|tma|all] > gpurun_out/sweep.txt
Shapes are picked through the tuning environment variables, which the library reads at every process call."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from ol_dsp_b200 import MODE_EXACT, MODE_FAST, MODE_FMA, BiquadBank, c2_coefs, noise_device  # noqa: E402

PEAK = 6548.2
MODES = {"fast": MODE_FAST, "fma": MODE_FMA, "exact": MODE_EXACT}


def run(C, F, mode, variant, env, steps=5, buf=None):
    for k in ("OLB_BQ_SKEW_CFG", "OLB_BQ_SYS_CFG", "OLB_BQ_TMA_CFG"):
        os.environ.pop(k, None)
    os.environ.update(env)
    x, y = buf
    b = BiquadBank(C, 8)
    b.SetMode(MODES[mode])
    b.SetVariant(variant)
    b.SetCoeffs(c2_coefs(C))
    st = torch.cuda.current_stream()
    try:
        for _ in range(2):
            b.ProcessDevice(x.data_ptr(), y.data_ptr(), F, 0, st.cuda_stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(steps):
            b.ProcessDevice(x.data_ptr(), y.data_ptr(), F, 0, st.cuda_stream)
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        v = C * F / (ms * 1e-3)
        cyc = ms * 1e-3 * 1.965e9 / F
        print(f"C{C:6d} F{F:7d} {mode:5s} v{variant} {str(env):40s} {b.LastKernel():10s} {ms:8.3f} ms  {v:.3e} ch-samp/s  frac {v * 8 / 1e9 / PEAK:.3f}  "
              f"{cyc:6.1f} cyc/frame", flush=True)
    except Exception as ex:  # noqa: BLE001
        print(f"C{C} {mode} v{variant} {env} FAILED {ex!r}", flush=True)
    b.close()


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "skew"
    torch.cuda.set_device(0)
    n = 65536 * 24000
    x = torch.empty(n, dtype=torch.float32, device="cuda")
    y = torch.empty(n, dtype=torch.float32, device="cuda")
    noise_device(x.data_ptr(), 1, 0, 65536, 0, 24000)
    buf = (x, y)
    shapes = {4096: (192000, ["48x32x4", "48x32x6", "96x16x4", "96x16x6", "48x16x4"]),
              8192: (96000, ["48x16x4", "48x16x6", "48x32x4", "96x16x4", "48x8x4"]),
              16384: (48000, ["48x8x4", "48x16x4", "24x8x4"]),
              32768: (48000, ["24x8x4", "48x8x4", "24x6x4"]),
              65536: (24000, ["24x4x4", "24x6x4", "24x8x4"])}
    for C, (F, cfgs) in shapes.items():
        if which in ("skew", "all"):
            for mode in ("fast", "fma", "exact"):
                for cfg in (cfgs if mode == "fast" else cfgs[:2]):
                    run(C, F, mode, 4, {"OLB_BQ_SKEW_CFG": cfg}, buf=buf)
        if which in ("tma", "all"):
            for mode in ("fast", "fma", "exact"):
                run(C, F, mode, 0, {}, buf=buf)
                run(C, F, mode, 2, {}, buf=buf)
            run(C, F, "fast", 3, {"OLB_BQ_SYS_CFG": "2x32x24x2"}, buf=buf)
