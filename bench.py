#!/usr/bin/env python
"""bench.py -- headline benchmark: channel-samples/s of the 8-section biquad cascade bank.

    python bench.py --gpus N --steps K --warmup W          (N>1: launched by torch.distributed.run)
    python bench.py --impl reference ...                   the CPU arm (oracle on the host cores)

Workload (config.workload): BASELINE.json configs[2] per GPU -- a 65,536-channel x 8-section cascade
bank at 48 kHz fed replayable white noise, streamed in blocks of --block-frames frames with the
recurrence state resident on the device between blocks (a "step" = one block of every channel).
Channels are independent, so N GPUs hold N disjoint 65,536-channel shards (weak scaling, no collective).
`--channels 4096` gives BASELINE.json configs[1] (C2).  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BYTES_PER_CHANNEL_SAMPLE = 8  # 4 B read + 4 B written per channel-sample
SEED = 0x01D5B200


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=100)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="olbank", choices=["olbank", "reference"])
    p.add_argument("--channels", type=int, default=65536)
    p.add_argument("--sections", type=int, default=8)
    p.add_argument("--block-frames", type=int, default=24000, help="frames per step (0.5 s of audio at 48 kHz)")
    p.add_argument("--mode", default="fast", choices=["fast", "fma", "exact"],
                   help="fast: normalised 4-op sections (throughput tier); fma: fused 5-op; exact: separately rounded (bit-exact tier)")
    p.add_argument("--variant", type=int, default=0)
    p.add_argument("--e2e-frames", type=int, default=4800)
    p.add_argument("--e2e-steps", type=int, default=10)
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-small", action="store_true", help="skip the 4096-channel configuration reported beside the headline")
    return p.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpu_arm(args, target_seconds):
    """Times the oracle's scalar cascade on the host cores: one channel per std::thread at a time,
    all hardware threads, on a bounded sample of the same workload (same coefficients, same noise)."""
    from oracle import loader
    from ol_dsp_b200 import c2_coefs
    orc = loader.oracle()
    threads = os.cpu_count() or 1
    mode = {"fast": loader.MODE_FAST, "fma": loader.MODE_FMA, "exact": loader.MODE_EXACT}[args.mode]
    frames = 48000
    probe_ch = threads
    k = c2_coefs(max(probe_ch, 64))[:, :args.sections]
    dt, _ = orc.biquad_bench(k[:probe_ch], None, frames, SEED, threads, mode)
    rate = probe_ch * frames / max(dt, 1e-6)
    channels = int(max(threads, min(args.channels, (rate * target_seconds / frames) // threads * threads)))
    k = c2_coefs(channels)[:, :args.sections]
    return orc, k, channels, frames, threads, mode


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    orc, k, channels, frames, threads, mode = cpu_arm(args, target_seconds=3.0)
    steps = max(1, min(args.steps, 5))
    warm = max(0, min(args.warmup, 1))
    for _ in range(warm):
        orc.biquad_bench(k, None, frames, SEED, threads, mode)
    total = 0.0
    for _ in range(steps):
        dt, _ = orc.biquad_bench(k, None, frames, SEED, threads, mode)
        total += dt
    value = channels * frames * steps / total
    sample = f"{channels} channels x {frames} frames per step ({steps} timed steps), same coefficients and Philox noise"
    line = {
        "impl": "reference", "metric": "channel-samples/s, 8-biquad cascade bank", "value": value, "unit": "channel-samples/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": 1e3 * total / steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": value, "unit": "channel-samples/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "channel-samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference has no biquad class (SURVEY.md 0.2): this arm is the scalar oracle cascade, one channel per "
                "std::thread at a time, all host threads; noise generation is inside the timed region",
    }
    print(json.dumps(line), flush=True)


def workload_config(args, n):
    return {
        "workload": f"{args.channels}-channel x {args.sections}-section biquad cascade bank per GPU, fp32 white noise (Philox), "
                    f"48 kHz, streamed in {args.block_frames}-frame blocks with device-resident state"
                    + (" [BASELINE configs[2] per GPU]" if args.channels == 65536 else " [BASELINE configs[1]]" if args.channels == 4096 else ""),
        "channels_per_gpu": args.channels, "sections": args.sections, "block_frames": args.block_frames, "sample_rate": 48000,
        "mode": args.mode, "layout": "[frames][channels]", "sharding": f"{n} x independent channel shards, no collective",
        "l2": f"no flush needed: each step streams {2 * 4 * args.channels * args.block_frames / 1e6:.0f} MB (in+out) >> 126 MB L2",
    }


def run_olbank(args):
    import ctypes as C

    import torch
    import torch.distributed as dist

    from ol_dsp_b200 import MODE_EXACT, MODE_FAST, MODE_FMA, BiquadBank, _lib, c2_coefs, noise_device, shard_range

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU: libolbank has no CPU fallback"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()

    Cn, S, F = args.channels, args.sections, args.block_frames
    bank = BiquadBank(Cn, S, 48000.0, device=local)
    bank.SetMode({"fast": MODE_FAST, "fma": MODE_FMA, "exact": MODE_EXACT}[args.mode])
    bank.SetVariant(args.variant)
    bank.SetCoeffs(c2_coefs(Cn)[:, :S])
    x = torch.empty((F, Cn), dtype=torch.float32, device="cuda")
    y = torch.empty_like(x)
    stream = torch.cuda.current_stream()
    first, _ = shard_range(world * Cn, rank, world)  # rank r owns channels [r*Cn, (r+1)*Cn) of the job
    noise_device(x.data_ptr(), SEED, first, Cn, 0, F, 1.0, 0, stream.cuda_stream)

    def step():
        bank.ProcessDevice(x.data_ptr(), y.data_ptr(), F, 0, stream.cuda_stream)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    launches = _lib.launch_count() - l0
    sampler.stop_flag = True
    sampler.join()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    kernel = bank.LastKernel()
    value = world * Cn * F * args.steps / (total_ms * 1e-3)

    # ---- e2e: host buffers through olb_biquad_process_host (pinned memory, copies inside the timed region)
    e2e = None
    if not args.no_e2e:
        EF = args.e2e_frames
        hin = torch.empty((EF, Cn), dtype=torch.float32).pin_memory()
        hout = torch.empty((EF, Cn), dtype=torch.float32).pin_memory()
        hin.copy_(x[:EF].cpu())
        bank.ProcessHostPtr(hin.data_ptr(), hout.data_ptr(), EF)  # warm-up (allocates staging)
        bank.ProcessHostPtr(hin.data_ptr(), hout.data_ptr(), EF)
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            bank.ProcessHostPtr(hin.data_ptr(), hout.data_ptr(), EF)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        barrier()
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": world * Cn * EF * args.e2e_steps / float(dt.item()), "unit": "channel-samples/s",
               "h2d_bytes_per_step": 4 * Cn * EF, "d2h_bytes_per_step": 4 * Cn * EF, "steps": args.e2e_steps,
               "frames_per_step": EF, "api": "olb_biquad_process_host (pinned host buffers, chunked H2D/kernel/D2H overlap)",
               "checksum": float(hout.double().sum().item())}

    # ---- BASELINE configs[1] beside the headline (N = 1 only): the 4 096-channel cascade, 10 s of audio as 192 000-frame blocks
    small = None
    if world == 1 and not args.no_small and Cn != 4096:
        C2, F2, K2 = 4096, 192000, 10
        if e2e:
            del hin, hout
        b2 = BiquadBank(C2, S, 48000.0, device=local)
        b2.SetMode({"fast": MODE_FAST, "fma": MODE_FMA, "exact": MODE_EXACT}[args.mode])
        b2.SetCoeffs(c2_coefs(C2)[:, :S])
        x2 = torch.empty((F2, C2), dtype=torch.float32, device="cuda")
        y2 = torch.empty_like(x2)
        noise_device(x2.data_ptr(), SEED, 0, C2, 0, F2, 1.0, 0, stream.cuda_stream)
        for _ in range(3):
            b2.ProcessDevice(x2.data_ptr(), y2.data_ptr(), F2, 0, stream.cuda_stream)
        torch.cuda.synchronize()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record(stream)
        for _ in range(K2):
            b2.ProcessDevice(x2.data_ptr(), y2.data_ptr(), F2, 0, stream.cuda_stream)
        s1.record(stream)
        torch.cuda.synchronize()
        ms2 = s0.elapsed_time(s1) / K2
        small = {"workload": "BASELINE configs[1]: 4096-channel x %d-section cascade, 48 kHz, streamed in 192000-frame blocks (4 s)" % S,
                 "value": C2 * F2 / (ms2 * 1e-3), "unit": "channel-samples/s", "ms_per_step": ms2, "steps": K2, "kernel": b2.LastKernel(),
                 "gbs": BYTES_PER_CHANNEL_SAMPLE * C2 * F2 / (ms2 * 1e-3) / 1e9}
        b2.close()
        del x2, y2

    if rank == 0:
        peak, peak_src = measured_peaks()
        per_launch_s = total_ms * 1e-3 / args.steps
        achieved = BYTES_PER_CHANNEL_SAMPLE * Cn * F / per_launch_s / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                traffic = json.load(open(tp)).get(kernel, {}).get(f"{Cn}x{F}")
            except Exception:
                traffic = None
        line = {
            "metric": "channel-samples/s, 8-biquad cascade bank", "value": value, "unit": "channel-samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args, world),
            "kernel": kernel, "gpu_launches": int(launches), "clocks": sampler.result(),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": BYTES_PER_CHANNEL_SAMPLE * Cn * F,
                         "fp32_lane_ops_per_channel_sample": {"fast": 4 * S + 1, "fma": 5 * S, "exact": 9 * S}[args.mode]},
        }
        if e2e:
            line["e2e"] = e2e
        if small:
            small["roofline_frac"] = small["gbs"] / peak
            line["configs1"] = small
        if world == 1 and not args.no_cpu_baseline:
            orc, k, channels, frames, threads, mode = cpu_arm(args, target_seconds=8.0)
            dt, _ = orc.biquad_bench(k, None, frames, SEED, threads, mode)
            line["cpu_baseline"] = {"value": channels * frames / dt, "unit": "channel-samples/s", "cores": threads, "kind": "port",
                                    "sample": f"{channels} channels x {frames} frames, scalar oracle cascade, one channel per std::thread at a time"}
        print(json.dumps(line), flush=True)
    bank.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_olbank(a)
