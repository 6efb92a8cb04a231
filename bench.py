#!/usr/bin/env python
"""bench.py -- headline benchmark: channel-samples/s of the 8-section biquad cascade bank.

    python bench.py --gpus N --steps K --warmup W          (N>1: launched by torch.distributed.run)
    python bench.py --impl reference ...                   the CPU arm (oracle on the host cores)

Workload (config.workload): BASELINE.json configs[2] per GPU -- a 65,536-channel x 8-section cascade
bank at 48 kHz fed replayable white noise, streamed in blocks of --block-frames frames with the
recurrence state resident on the device between blocks (a "step" = one block of every channel).
Channels are independent, so N GPUs hold N disjoint 65,536-channel shards (weak scaling, no collective).
`--channels 4096` gives BASELINE.json configs[1] (C2).  Prints ONE JSON line on rank 0.

Beside the headline (`value`, device-resident, FAST tier) the line carries, all measured in the same run:
  tiers         the same launch in the three arithmetic tiers (EXACT is the bit-exact one)
  parity_check  one more step of the headline launch, checked bit for bit against the oracle for 256 channels
  sustained     >= 2 s of back-to-back headline launches with the NVML clock / reason sample
  strong        the SAME 65,536-channel bank cut into N shards (BASELINE configs[2] as written): 65,536 / N channels per GPU
  voices        BASELINE configs[3]: 65,536 voices per GPU, the stereo bus reduced across the GPUs by the library's NCCL
                all-reduce on its own stream, under the next block's voices
  configs1      BASELINE configs[1] (4,096 channels), N = 1
  e2e           the metric through olb_biquad_process_host (pinned HOST buffers near the GPU's NUMA node, H2D and D2H inside
                the timed region), with the box's concurrent copy ceiling measured in the same run (pcie_frac)
  cpu_baseline  the oracle's scalar cascade on the host cores, N = 1
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BYTES_PER_CHANNEL_SAMPLE = 8  # 4 B read + 4 B written per channel-sample
SEED = 0x01D5B200
LANE_OPS = {"fast": lambda s: 4 * s + 1, "fma": lambda s: 5 * s, "exact": lambda s: 9 * s}


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
    p.add_argument("--e2e-seconds", type=float, default=2.0)
    p.add_argument("--sustain-seconds", type=float, default=2.0)
    p.add_argument("--voices", type=int, default=65536, help="voices per GPU of the `voices` record")
    p.add_argument("--voice-frames", type=int, default=4800)
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-small", action="store_true", help="skip the 4096-channel configuration reported beside the headline")
    p.add_argument("--no-tiers", action="store_true")
    p.add_argument("--no-parity", action="store_true")
    p.add_argument("--no-sustained", action="store_true")
    p.add_argument("--no-strong", action="store_true")
    p.add_argument("--no-voices", action="store_true")
    p.add_argument("--only-headline", action="store_true", help="tuning aid: value + roofline only")
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
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz, self.power = index, [], set(), False, None, []
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
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def finish(self):
        self.stop_flag = True
        self.join()
        return self.result()

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "power_w_max": max(self.power) if self.power else None}


def git_head():
    try:
        return subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True, timeout=5).stdout.strip() or None
    except Exception:
        return None


# ---- the C2 / C3 coefficient workload, from the ORACLE's designer (reference arm and cpu_baseline: nothing of libolbank) ----
def c2_coefs_oracle(orc, channels, sample_rate=48000.0):
    """Same table as ol_dsp_b200.workloads.c2_coefs (tests/test_abi.py holds the two designers bit-equal)."""
    import numpy as np
    base = np.empty((64, 8, 5), dtype=np.float32)

    def rbj(typ, f0, q, gain_db=0.0):
        out = np.empty(5, dtype=np.float32)
        orc.lib.orc_rbj(typ, sample_rate, f0, q, gain_db, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float)))
        return out

    for m in range(64):
        fc = 200.0 * (12000.0 / 200.0) ** (m / 63.0)
        for s in range(4):
            base[m, s] = rbj(0, fc, 0.707)
        base[m, 4] = rbj(2, min(fc * 1.5, 0.45 * sample_rate), 2.0, 3.0)
        base[m, 5] = rbj(2, max(fc * 0.5, 40.0), 2.0, -3.0)
        base[m, 6] = rbj(1, 30.0, 0.707)
        base[m, 7] = rbj(3, min(fc * 2.0, 0.45 * sample_rate), 1.0)
    return np.ascontiguousarray(base[np.arange(channels) % 64])


def cpu_arm(args, target_seconds):
    """The oracle's scalar cascade on the host cores: one channel per std::thread at a time, all hardware threads, on a
    bounded sample of the same workload (same coefficients, same noise).  -> (orc, coefs, channels, frames, threads, mode)"""
    from oracle import loader
    orc = loader.oracle()
    threads = os.cpu_count() or 1
    mode = {"fast": loader.MODE_FAST, "fma": loader.MODE_FMA, "exact": loader.MODE_EXACT}[args.mode]
    frames = 48000
    probe_ch = threads
    k = c2_coefs_oracle(orc, max(probe_ch, 64))[:, :args.sections]
    dt, _ = orc.biquad_bench(k[:probe_ch], None, frames, SEED, threads, mode)
    rate = probe_ch * frames / max(dt, 1e-6)
    channels = int(max(threads, min(args.channels, (rate * target_seconds / frames) // threads * threads)))
    k = c2_coefs_oracle(orc, channels)[:, :args.sections]
    return orc, k, channels, frames, threads, mode


CPU_NOTE = ("scalar oracle cascade (the reference has no biquad class, SURVEY.md 0.2), one channel per std::thread at a time, not "
            "vectorised across channels; noise generation and the checksum pass are outside the timed region")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # each step is a bounded sample of the workload sized so that --steps of them end within a few minutes
    per_step = max(0.5, min(3.0, 150.0 / max(1, args.steps + args.warmup)))
    orc, k, channels, frames, threads, mode = cpu_arm(args, target_seconds=per_step)
    steps, warm = max(1, args.steps), max(0, args.warmup)
    total, _ = orc.biquad_bench(k, None, frames, SEED, threads, mode, steps=steps, warmup=warm)  # one input set, W + K passes over it
    value = channels * frames * steps / total
    sample = f"{channels} channels x {frames} frames per step ({steps} timed steps, {warm} warm-up), same coefficients and Philox noise"
    line = {
        "impl": "reference", "metric": "channel-samples/s, 8-biquad cascade bank", "value": value, "unit": "channel-samples/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": 1e3 * total / steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": value, "unit": "channel-samples/s", "cores": threads, "kind": "port", "sample": sample, "note": CPU_NOTE},
        "e2e": {"value": value, "unit": "channel-samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


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
    import numpy as np
    import torch
    import torch.distributed as dist

    from ol_dsp_b200 import MODE_EXACT, MODE_FAST, MODE_FMA, BiquadBank, Comm, VoiceBank, _lib, c2_coefs, noise_device, shard_range

    MODES = {"fast": MODE_FAST, "fma": MODE_FMA, "exact": MODE_EXACT}
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU: libolbank has no CPU fallback"
    torch.cuda.set_device(local)
    L = _lib.lib()
    numa = ctypes.c_int(-1)
    _lib.check(L.olb_bind_thread_near(local, ctypes.byref(numa)))  # this rank's host thread next to its GPU, before anything is allocated
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peak, peak_src = measured_peaks()
    if args.only_headline:
        args.no_tiers = args.no_parity = args.no_sustained = args.no_strong = args.no_voices = args.no_small = args.no_e2e = args.no_cpu_baseline = True

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(v):
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    stream = torch.cuda.current_stream()
    sid = stream.cuda_stream

    def timed(fn, steps, warm=3, clocks=False):
        """-> (ms per step, max over ranks; clock sample or None).  CUDA events on the launch stream, barrier on both sides."""
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        barrier()
        sampler = ClockSampler(local) if clocks else None
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        ck = sampler.finish() if sampler else None
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)) / steps, ck

    Cn, S, F = args.channels, args.sections, args.block_frames
    coefs = c2_coefs(Cn)[:, :S]
    bank = BiquadBank(Cn, S, 48000.0, device=local)
    bank.SetMode(MODES[args.mode])
    bank.SetVariant(args.variant)
    bank.SetCoeffs(coefs)
    x = torch.empty((F, Cn), dtype=torch.float32, device="cuda")
    y = torch.empty_like(x)
    first, _ = shard_range(world * Cn, rank, world)  # rank r owns channels [r*Cn, (r+1)*Cn) of the job
    noise_device(x.data_ptr(), SEED, first, Cn, 0, F, 1.0, 0, sid)

    def step():
        bank.ProcessDevice(x.data_ptr(), y.data_ptr(), F, 0, sid)

    def frac_of(cs_per_s):
        return BYTES_PER_CHANNEL_SAMPLE * cs_per_s / 1e9 / peak

    # ---- headline ------------------------------------------------------------------------------------------------
    warm = max(args.warmup, 3)
    l0 = _lib.launch_count()
    ms, clocks = timed(step, args.steps, warm, clocks=True)
    launches = (_lib.launch_count() - l0) * args.steps // (args.steps + warm)
    kernel = bank.LastKernel()
    value = world * Cn * F / (ms * 1e-3)

    # ---- parity_check: one more step of the same launch, from the bank's read-back state, against the oracle ---------
    parity = None
    if not args.no_parity and rank == 0:
        from oracle import loader
        orc = loader.oracle()
        omode = {"fast": loader.MODE_FAST, "fma": loader.MODE_FMA, "exact": loader.MODE_EXACT}[args.mode]
        pick = np.unique(np.r_[0:64, Cn // 3:Cn // 3 + 64, Cn // 2:Cn // 2 + 32, np.linspace(0, Cn - 1, 64).astype(np.int64), Cn - 64:Cn] % Cn)
        st = bank.GetState()[pick]
        step()
        torch.cuda.synchronize()
        idx = torch.as_tensor(pick, device="cuda")
        got, xin = y[:, idx].cpu().numpy(), x[:, idx].cpu().numpy()
        want = orc.biquad(xin, np.ascontiguousarray(coefs[pick]), mode=omode, state=np.ascontiguousarray(st))
        ok = bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)))
        parity = {"channels": int(len(pick)), "frames": F, "mode": args.mode, "kernel": bank.LastKernel(), "bit_exact": ok,
                  "how": "one more step of the timed launch (same buffers, state read back from the bank before it); oracle fed the same "
                         "input block and state; channels from the first, middle and last warps of the grid"}
    if world > 1:
        barrier()

    # ---- sustained: >= sustain_seconds of back-to-back headline launches -----------------------------------------
    sustained = None
    if not args.no_sustained:
        n = max(args.steps, int(args.sustain_seconds / (ms * 1e-3)) + 1)
        ms_s, ck_s = timed(step, n, 1, clocks=True)
        v = world * Cn * F / (ms_s * 1e-3)
        sustained = {"seconds": n * ms_s * 1e-3, "steps": n, "value": v, "ms_per_step": ms_s, "roofline_frac": frac_of(v / world), "clocks": ck_s}

    # ---- tiers: the same launch in the three arithmetic tiers ------------------------------------------------------
    tiers = None
    if not args.no_tiers:
        tiers = {}
        for name in ("exact", "fma", "fast"):
            bank.SetMode(MODES[name])
            bank.Init(48000.0)
            ms_t, _ = timed(step, max(10, args.steps // 5), 3)
            v = world * Cn * F / (ms_t * 1e-3)
            tiers[name] = {"value": v, "ms_per_step": ms_t, "roofline_frac": frac_of(v / world), "kernel": bank.LastKernel(),
                           "fp32_lane_ops_per_channel_sample": LANE_OPS[name](S),
                           "parity": "bit-exact vs the oracle's separately rounded cascade" if name == "exact" else
                                     "bit-exact vs the oracle's restatement of the same fused operations; tolerance tier vs EXACT (BASELINE.md)"}
        bank.SetMode(MODES[args.mode])
        bank.Init(48000.0)

    # ---- strong: the SAME bank cut into `world` shards (BASELINE configs[2] as written) ---------------------------
    strong = None
    if not args.no_strong:
        if world == 1:
            strong = {"channels_total": Cn, "channels_per_gpu": Cn, "value": value, "ms_per_step": ms, "kernel": kernel,
                      "roofline_frac": frac_of(value), "note": "N = 1: the headline launch"}
        else:
            Cs = Cn // world
            Fs = F * world  # the same bytes per GPU and step as the headline
            f0, _ = shard_range(Cn, rank, world)
            bs = BiquadBank(Cs, S, 48000.0, device=local)
            bs.SetMode(MODES[args.mode])
            bs.SetCoeffs(np.ascontiguousarray(c2_coefs(Cn)[f0:f0 + Cs, :S]))
            xs, ys = x.view(-1)[:Fs * Cs].view(Fs, Cs), y.view(-1)[:Fs * Cs].view(Fs, Cs)
            noise_device(xs.data_ptr(), SEED, f0, Cs, 0, Fs, 1.0, 0, sid)
            ms_g, _ = timed(lambda: bs.ProcessDevice(xs.data_ptr(), ys.data_ptr(), Fs, 0, sid), max(10, args.steps // 5), 3)
            v = Cn * Fs / (ms_g * 1e-3)
            strong = {"channels_total": Cn, "channels_per_gpu": Cs, "block_frames": Fs, "value": v, "ms_per_step": ms_g,
                      "kernel": bs.LastKernel(), "roofline_frac": frac_of(v / world),
                      "note": "total channel-samples/s of ONE 65,536-channel bank sharded by channel range; divide by the N = 1 line's value for the strong-scaling factor"}
            bs.close()
            noise_device(x.data_ptr(), SEED, first, Cn, 0, F, 1.0, 0, sid)

    # ---- voices: BASELINE configs[3] -------------------------------------------------------------------------------
    voices = None
    if not args.no_voices:
        V, VF = args.voices, args.voice_frames
        vb = VoiceBank(V, device=local)
        for cc, val in ((41, 70), (42, 30), (73, 64), (75, 60), (108, 10), (111, 40)):  # presets through the MIDI maps (main.cpp:114-128)
            vb.UpdateMidiControl(cc, val)
        vb.NoteOn((36 + (rank * V + np.arange(V)) % 61).astype(np.uint8))
        comm = Comm.from_torch_distributed(local) if world > 1 else None
        buses = [torch.zeros((VF, 2), dtype=torch.float32, device="cuda") for _ in range(2)]
        k = [0]

        def vstep():
            slot = k[0] % 2  # block k -> bus k % 2; its reduce runs on the communicator's stream under block k + 1's voices
            if comm:
                comm.Wait(slot, sid)
            vb.ProcessDevice(VF, bus_ptr=buses[slot].data_ptr(), bus_channels=2, bus_gain=1.0 / (V * world), stream=sid)
            if comm:
                comm.AllReduceBusAsync(buses[slot].data_ptr(), VF, 2, sid, slot)
            k[0] += 1

        def vsteps_done():
            if comm:
                comm.Wait(0, sid)
                comm.Wait(1, sid)

        for _ in range(3):
            vstep()
        vsteps_done()
        torch.cuda.synchronize()
        barrier()
        nv = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(nv):
            vstep()
        vsteps_done()
        e1.record(stream)
        torch.cuda.synchronize()
        barrier()
        ms_v = max_over_ranks(e0.elapsed_time(e1)) / nv
        voices = {"workload": f"BASELINE configs[3]: {V} SynthVoice per GPU (PolyBLEP saw + 2 ADSR + portamento + SVF), note 36 + (v mod 61), stereo bus",
                  "voices_per_gpu": V, "frames_per_step": VF, "steps": nv, "value": world * V * VF / (ms_v * 1e-3), "unit": "voice-samples/s",
                  "ms_per_step": ms_v, "kernel": "k_voice + k_bus_reduce",
                  "collective": ("olb_voice_bus_allreduce_async (libolbank.so: one NCCL fp32 sum all-reduce of the [frames][2] bus per step, on the "
                                 "communicator's stream under the next step's voices)") if comm else "none (1 GPU)",
                  "bus_bytes_per_step": 8 * VF, "bus_checksum": float(buses[(k[0] - 1) % 2].double().sum().item())}
        if comm:
            comm.close()
        vb.close()

    # ---- BASELINE configs[1] beside the headline (N = 1 only): the 4 096-channel cascade, 10 s of audio as 192 000-frame blocks
    small = None
    if world == 1 and not args.no_small and Cn != 4096:
        C2, F2 = 4096, 192000
        b2 = BiquadBank(C2, S, 48000.0, device=local)
        b2.SetMode(MODES[args.mode])
        b2.SetCoeffs(c2_coefs(C2)[:, :S])
        x2, y2 = x.view(-1)[:F2 * C2].view(F2, C2), y.view(-1)[:F2 * C2].view(F2, C2)
        noise_device(x2.data_ptr(), SEED, 0, C2, 0, F2, 1.0, 0, sid)
        ms2, _ = timed(lambda: b2.ProcessDevice(x2.data_ptr(), y2.data_ptr(), F2, 0, sid), 10, 3)
        small = {"workload": "BASELINE configs[1]: 4096-channel x %d-section cascade, 48 kHz, streamed in 192000-frame blocks (4 s)" % S,
                 "value": C2 * F2 / (ms2 * 1e-3), "unit": "channel-samples/s", "ms_per_step": ms2, "steps": 10, "kernel": b2.LastKernel(),
                 "gbs": BYTES_PER_CHANNEL_SAMPLE * C2 * F2 / (ms2 * 1e-3) / 1e9}
        small["roofline_frac"] = small["gbs"] / peak
        b2.close()
        noise_device(x.data_ptr(), SEED, first, Cn, 0, F, 1.0, 0, sid)

    # ---- e2e: host buffers through olb_biquad_process_host (pinned, near the GPU's NUMA node; copies inside the timed region)
    e2e = None
    if not args.no_e2e:
        EF = args.e2e_frames
        nbytes = 4 * Cn * EF
        hp_in, hp_out = ctypes.c_void_p(), ctypes.c_void_p()
        _lib.check(L.olb_host_alloc_near(local, ctypes.byref(hp_in), nbytes, None))
        _lib.check(L.olb_host_alloc_near(local, ctypes.byref(hp_out), nbytes, None))
        hin = np.ctypeslib.as_array(ctypes.cast(hp_in, ctypes.POINTER(ctypes.c_float)), shape=(EF, Cn))
        hout = np.ctypeslib.as_array(ctypes.cast(hp_out, ctypes.POINTER(ctypes.c_float)), shape=(EF, Cn))
        hin[:] = x[:EF].cpu().numpy()
        hout[:] = 0

        # (1) what the box gives N ranks copying both ways at once: plain cudaMemcpyAsync of the same buffers, 64 MiB chunks
        s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
        hin_t, hout_t = torch.from_numpy(hin), torch.from_numpy(hout)  # cudaHostAlloc memory: torch sees it as pinned
        assert hin_t.is_pinned() and hout_t.is_pinned()
        rows = max(1, (64 << 20) // (4 * Cn))

        def raw_copies():
            with torch.cuda.stream(s_up):
                for r in range(0, EF, rows):
                    n = min(rows, EF - r)
                    x[r:r + n].copy_(hin_t[r:r + n], non_blocking=True)    # H2D
            with torch.cuda.stream(s_dn):
                for r in range(0, EF, rows):
                    n = min(rows, EF - r)
                    hout_t[r:r + n].copy_(y[r:r + n], non_blocking=True)   # D2H

        def wall(fn, reps):
            torch.cuda.synchronize()
            barrier()
            t0 = time.perf_counter()
            for _ in range(reps):
                fn()
            torch.cuda.synchronize()
            dt = max_over_ranks(time.perf_counter() - t0)
            barrier()
            return dt

        raw_copies()
        t1 = wall(raw_copies, 2) / 2
        raw_reps = max(3, int(1.0 / max(t1, 1e-4)))
        t_raw = wall(raw_copies, raw_reps) / raw_reps
        ceiling = nbytes / t_raw / 1e9  # GB/s per GPU, each direction, all ranks copying at once
        noise_device(x.data_ptr(), SEED, first, Cn, 0, F, 1.0, 0, sid)
        torch.cuda.synchronize()

        # (2) the product path
        def host_step():
            bank.ProcessHostPtr(hp_in.value, hp_out.value, EF)

        host_step()
        host_step()
        t1 = wall(host_step, 2) / 2
        reps = max(5, int(args.e2e_seconds / max(t1, 1e-4)) + 1)
        dt = wall(host_step, reps)
        e2e_gbs = nbytes * reps / dt / 1e9
        e2e = {"value": world * Cn * EF * reps / dt, "unit": "channel-samples/s",
               "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes, "steps": reps, "seconds": dt,
               "frames_per_step": EF, "api": "olb_biquad_process_host (pinned host buffers from olb_host_alloc_near, chunked H2D / kernel / D2H overlap)",
               "gbs_each_way_per_gpu": e2e_gbs, "pcie_ceiling_gbs_each_way_per_gpu": ceiling, "pcie_frac": e2e_gbs / ceiling,
               "pcie_ceiling_how": f"{world} rank(s) at once: plain async H2D + D2H copies of the same pinned buffers in 64 MiB chunks on two streams, "
                                   f"{raw_reps} repetitions, max over ranks",
               "numa_node": int(numa.value), "checksum": float(hout.astype(np.float64).sum())}
        del hin, hout, hin_t, hout_t
        L.olb_host_free(hp_in)
        L.olb_host_free(hp_out)

    if rank == 0:
        per_launch_s = ms * 1e-3
        achieved = BYTES_PER_CHANNEL_SAMPLE * Cn * F / per_launch_s / 1e9
        traffic, traffic_commit = None, None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                tj = json.load(open(tp))
                traffic = tj.get(kernel, {}).get(f"{Cn}x{F}")
                traffic_commit = tj.get("commit")
            except Exception:
                traffic = None
        line = {
            "metric": "channel-samples/s, 8-biquad cascade bank", "value": value, "unit": "channel-samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": warm, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args, world),
            "kernel": kernel, "gpu_launches": int(launches), "clocks": clocks, "head": git_head(),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_commit": traffic_commit,
                         "traffic_how": "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum of one launch of this kernel at this shape, captured at traffic_commit (profiles/traffic.json)",
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": BYTES_PER_CHANNEL_SAMPLE * Cn * F,
                         "fp32_lane_ops_per_channel_sample": LANE_OPS[args.mode](S)},
        }
        for key, val in (("parity_check", parity), ("sustained", sustained), ("tiers", tiers), ("strong", strong), ("voices", voices),
                         ("e2e", e2e), ("configs1", small)):
            if val is not None:
                line[key] = val
        if world == 1 and not args.no_cpu_baseline:
            orc, k, channels, frames, threads, mode = cpu_arm(args, target_seconds=8.0)
            dt, _ = orc.biquad_bench(k, None, frames, SEED, threads, mode)
            line["cpu_baseline"] = {"value": channels * frames / dt, "unit": "channel-samples/s", "cores": threads, "kind": "port",
                                    "sample": f"{channels} channels x {frames} frames", "note": CPU_NOTE}
        emit(line)
        if parity is not None and not parity["bit_exact"]:
            print("bench.py: parity_check FAILED: the timed launch does not match the oracle", file=sys.stderr)
            bank.close()
            sys.exit(3)
    bank.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def emit(line):
    """The ONE JSON line, on the process's real stdout."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


if __name__ == "__main__":
    a = parse()
    # stdout carries exactly one JSON line: whatever libraries print there meanwhile (NCCL's version banner, build chatter)
    # goes to stderr
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_olbank(a)
