"""Development aid: does running part of a wide bank as whole-cascade warps (k_bq_tma) and the rest as fine-grained units
(k_bq_skew) CONCURRENTLY balance the 592 schedulers better than k_bq_tma alone (1 024 warps = 1.73 per scheduler)?
Two banks on two streams, each with its own buffers; total channel-samples/s against one 65 536-channel bank.
    python tools/hybrid_probe.py [frames]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from ol_dsp_b200 import MODE_FAST, BiquadBank, c2_coefs, noise_device  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 24000
TOTAL = 65536


def bank(C, variant):
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FAST)
    b.SetVariant(variant)
    b.SetCoeffs(c2_coefs(C))
    x = torch.empty((F, C), dtype=torch.float32, device="cuda")
    noise_device(x.data_ptr(), 1, 0, C, 0, F)
    return b, x, torch.empty_like(x)


def timed(parts, steps=5):
    streams = [torch.cuda.Stream() for _ in parts]
    def go():
        for (b, x, y), s in zip(parts, streams):
            b.ProcessDevice(x.data_ptr(), y.data_ptr(), F, 0, s.cuda_stream)
    for _ in range(2):
        go()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    main = torch.cuda.current_stream()
    e0.record(main)
    for s in streams:
        s.wait_stream(main)
    for _ in range(steps):
        go()
    for s in streams:
        main.wait_stream(s)
    e1.record(main)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


one = bank(TOTAL, 0)
ms = timed([one])
print(f"one bank 65536 {one[0].LastKernel()} {ms:.3f} ms {TOTAL * F / ms / 1e-3:.3e}", flush=True)
one[0].close()
del one
for whole in (37888, 28416, 47360, 18944, 56832):
    for cfg in ("24x8x4", "24x4x4", "48x8x4"):
        os.environ["OLB_BQ_SKEW_CFG"] = cfg
        a, b = bank(whole, 2), bank(TOTAL - whole, 4)
        ms = timed([a, b])
        print(f"tma {whole} + skew {TOTAL - whole} ({cfg}) {ms:.3f} ms {TOTAL * F / ms / 1e-3:.3e}  frac {TOTAL * F * 8 / ms / 1e-3 / 1e9 / 6548.2:.3f}", flush=True)
        a[0].close(); b[0].close()
        del a, b
