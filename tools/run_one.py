"""Development aid: one bank shape, a few launches (the command ncu wraps).
    python tools/run_one.py <channels> <frames> <mode> <variant> [steps] [sections]      env: OLB_BQ_*_CFG"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from ol_dsp_b200 import MODE_EXACT, MODE_FAST, MODE_FMA, BiquadBank, c2_coefs, noise_device  # noqa: E402

C, F, mode, variant = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4])
steps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
S = int(sys.argv[6]) if len(sys.argv) > 6 else 8
x = torch.empty((F, C), dtype=torch.float32, device="cuda")
y = torch.empty_like(x)
noise_device(x.data_ptr(), 1, 0, C, 0, F)
b = BiquadBank(C, S)
b.SetMode({"fast": MODE_FAST, "fma": MODE_FMA, "exact": MODE_EXACT}[mode])
b.SetVariant(variant)
b.SetCoeffs(c2_coefs(C)[:, :S])
st = torch.cuda.current_stream()
b.ProcessDevice(x.data_ptr(), y.data_ptr(), F, 0, st.cuda_stream)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(steps):
    b.ProcessDevice(x.data_ptr(), y.data_ptr(), F, 0, st.cuda_stream)
e1.record(st)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
print(f"C{C} F{F} S{S} {mode} v{variant} {b.LastKernel()} {ms:.3f} ms {C * F / ms / 1e-3:.3e} ch-samp/s {ms * 1e-3 * 1.965e9 / F:.1f} cyc/frame", flush=True)
b.close()
