"""Secondary benchmarks (instance-samples/s for the non-headline rows).  One JSON line per unit.
    python tools/bench_units.py dattorro [--instances 8192] [--frames 48000] [--steps 3]
CUDA events on the launch stream; inputs resident in HBM; state (rings) persists across steps like a running plate."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return float(json.load(open(p))["hbm_gbs"]) if os.path.exists(p) else 6650.0


def timed(fn, steps, warm=2):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / steps


def bench_dattorro(a):
    import torch
    from ol_dsp_b200 import DattorroBank, noise_device
    I, F = a.instances, a.frames
    b = DattorroBank(I)
    b.SetDecay((0.5 + 0.4 * (torch.arange(I) % 16) / 16).numpy())
    x = torch.empty((F, I), dtype=torch.float32, device="cuda")
    noise_device(x.data_ptr(), 0x01D5B200, 0, I, 0, F, 0.25)
    l, r = torch.empty_like(x), torch.empty_like(x)
    st = torch.cuda.current_stream().cuda_stream
    dt = timed(lambda: b.ProcessDevice(x.data_ptr(), l.data_ptr(), r.data_ptr(), F, st), a.steps)
    units = I * F
    line = {"unit": "dattorro", "metric": "instance-samples/s", "value": units / dt, "ms_per_step": dt * 1e3, "instances": I, "frames": F,
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_unit": 12, "achieved": 12 * units / dt / 1e9, "peak": peaks(), "unit": "GB/s",
                         "frac": 12 * units / dt / 1e9 / peaks(),
                         "ring_bytes_per_unit": 4 * (13 + 13 + 14), "ring_traffic_gbs": 160 * units / dt / 1e9},
            "env": {k: v for k, v in os.environ.items() if k.startswith("OLB_")},
            "checksum": float(l.double().sum().item() + r.double().sum().item())}
    if a.cpu:
        import time

        import numpy as np
        from oracle import loader
        ref = loader.reference() or loader.oracle()
        xs = np.ascontiguousarray(x[:, 0].cpu().numpy())
        t0 = time.perf_counter()
        ref.dattorro(xs, None)
        line["cpu_baseline"] = {"value": F / (time.perf_counter() - t0), "unit": "instance-samples/s", "cores": 1,
                                "kind": "reference" if loader.reference() else "port", "sample": f"1 instance x {F} frames, verb.cpp compiled unmodified"}
    print(json.dumps(line), flush=True)
    b.close()


def bench_fx(a, kind, name, bytes_per_unit):
    import numpy as np
    import torch
    from ol_dsp_b200 import FxBank, noise_device
    I, F = a.instances, a.frames
    b = FxBank(kind, I)
    b.SetMode({"exact": 0, "fma": 1}[a.mode])
    b.UpdateHardwareControl(35, (0.05 + 0.9 * np.arange(I) / I).astype(np.float32))  # CC_DELAY_TIME sweep
    x = torch.empty((F, I), dtype=torch.float32, device="cuda")
    noise_device(x.data_ptr(), 0x01D5B200, 0, I, 0, F, 0.25)
    y = torch.empty_like(x)
    st = torch.cuda.current_stream().cuda_stream
    dt = timed(lambda: b.ProcessDevice(x.data_ptr(), y.data_ptr(), F, st), a.steps)
    units = I * F
    line = {"unit": name, "mode": a.mode, "metric": "instance-samples/s", "value": units / dt, "ms_per_step": dt * 1e3, "instances": I, "frames": F,
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_unit": bytes_per_unit, "achieved": bytes_per_unit * units / dt / 1e9,
                         "peak": peaks(), "unit": "GB/s", "frac": bytes_per_unit * units / dt / 1e9 / peaks()},
            "kernel": b.LastKernel(), "env": {k: v for k, v in os.environ.items() if k.startswith("OLB_")},
            "checksum": float(y.double().sum().item())}
    if a.cpu:
        import time
        from oracle import loader
        ref = loader.reference() or loader.oracle()
        xs = np.ascontiguousarray(x[:, :1].cpu().numpy())
        t0 = time.perf_counter()
        getattr(ref, name)(xs, 48000.0, [])
        line["cpu_baseline"] = {"value": F / (time.perf_counter() - t0), "unit": "instance-samples/s", "cores": 1,
                                "kind": "reference" if loader.reference() else "port",
                                "sample": f"1 instance x {F} frames, the reference's own {name} class over the DaisySP restatement"}
    print(json.dumps(line), flush=True)
    b.close()


def bench_chorus(a):
    """C5: chorus rate / depth swept per instance."""
    import numpy as np
    import torch
    from ol_dsp_b200 import ChorusBank, noise_device
    I, F = a.instances, a.frames
    b = ChorusBank(I)
    b.Set("rate", ((np.arange(I) % 64) / 63.0).astype(np.float32))
    b.Set("depth", (0.08 + 0.92 * np.arange(I) / I).astype(np.float32))
    b.Set("phase", ((np.arange(I) % 7) / 7.0).astype(np.float32))
    x = torch.empty((F, I), dtype=torch.float32, device="cuda")
    noise_device(x.data_ptr(), 0x01D5B200, 0, I, 0, F, 0.25)
    y = torch.empty_like(x)
    st = torch.cuda.current_stream().cuda_stream
    dt = timed(lambda: b.ProcessDevice(x.data_ptr(), y.data_ptr(), F, st), a.steps)
    units = I * F
    line = {"unit": "chorus", "metric": "instance-samples/s", "value": units / dt, "ms_per_step": dt * 1e3, "instances": I, "frames": F,
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_unit": 8, "achieved": 8 * units / dt / 1e9, "peak": peaks(), "unit": "GB/s",
                         "frac": 8 * units / dt / 1e9 / peaks()},
            "kernel": b.LastKernel(), "checksum": float(y.double().sum().item())}
    if a.cpu:
        import time
        from oracle import loader
        xs = np.ascontiguousarray(x[:, 0].cpu().numpy())
        t0 = time.perf_counter()
        loader.oracle().chorus(xs)
        line["cpu_baseline"] = {"value": F / (time.perf_counter() - t0), "unit": "instance-samples/s", "cores": 1, "kind": "port",
                                "sample": f"1 instance x {F} frames, the oracle's scalar chorus"}
    print(json.dumps(line), flush=True)
    b.close()


def bench_voices(a):
    """C4: `instances` voices PER GPU (weak), note 36 + (v mod 61), presets through the MIDI maps, NoteOn at t = 0; each rank
    reduces its voices to a partial stereo bus on chip, then ONE NCCL all-reduce of [frames][2] floats per step."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from ol_dsp_b200 import Comm, VoiceBank
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    V, F = a.instances, a.frames
    b = VoiceBank(V, device=local)
    for cc, val in ((41, 70), (42, 30), (73, 64), (75, 60), (108, 10), (111, 40)):
        b.UpdateMidiControl(cc, val)
    b.NoteOn((36 + (rank * V + np.arange(V)) % 61).astype(np.uint8))
    comm = Comm.from_torch_distributed(local) if world > 1 else None
    buses = [torch.zeros((F, 2), dtype=torch.float32, device="cuda") for _ in range(2)]
    bus = buses[0]
    st = torch.cuda.current_stream().cuda_stream
    k = [0]

    def step():
        # block k's voices into bus k % 2; its reduce runs on the communicator's stream under block k + 1's voices
        slot = k[0] % 2
        if comm:
            comm.Wait(slot, st)
        b.ProcessDevice(F, bus_ptr=buses[slot].data_ptr(), bus_channels=2, bus_gain=1.0 / (V * world), stream=st)
        if comm:
            comm.AllReduceBusAsync(buses[slot].data_ptr(), F, 2, st, slot)
        k[0] += 1

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    if comm:
        comm.Wait(0, st)
        comm.Wait(1, st)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    dt = float(ms.item()) * 1e-3 / a.steps
    if rank == 0:
        units = V * world * F
        print(json.dumps({"unit": "voices", "metric": "voice-samples/s", "value": units / dt, "ms_per_step": dt * 1e3, "n_gpus": world,
                          "voices_per_gpu": V, "frames": F, "scaling": "weak",
                          "collective": "olb_voice_bus_allreduce_async: one NCCL all-reduce of the [frames][2] fp32 bus per step, on the communicator's stream under the next step's voices" if world > 1 else "none (1 GPU)",
                          "bus_bytes_per_step": 8 * F, "bound": "fp32/sfu issue (no input stream; 8 B/frame/GPU out)",
                          "checksum": float(bus.double().sum().item())}), flush=True)
    b.close()
    if world > 1:
        dist.destroy_process_group()


def bench_sampler(a):
    """Sample playback bank: a kit of 128 one-second samples, every instance looping one of them
    from its own start offset (the pool is 24.6 MB: it lives in L2, the output stream goes to HBM)."""
    import numpy as np
    import torch
    from ol_dsp_b200 import SamplerBank, sampler
    I, F = a.instances, a.frames
    b = SamplerBank(I, 1)
    r = np.random.default_rng(7)
    ids = [b.LoadSample(r.standard_normal(48000).astype(np.float32)) for _ in range(128)]
    b.SetSample(np.array([ids[i % 128] for i in range(I)], dtype=np.uint32))
    b.SetPlayMode(sampler.LOOP)
    b.Play()
    for k in range(64):  # 64 start offsets across the bank
        b.Seek(741 * k, first=k * (I // 64), count=I // 64)
    y = torch.zeros((F, I), dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    dt = timed(lambda: b.ProcessDevice(y.data_ptr(), F, st), a.steps)
    units = I * F
    line = {"unit": "sampler", "metric": "instance-samples/s", "value": units / dt, "ms_per_step": dt * 1e3, "instances": I, "frames": F,
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_unit": 8, "achieved": 8 * units / dt / 1e9, "peak": peaks(), "unit": "GB/s",
                         "frac": 8 * units / dt / 1e9 / peaks()},
            "kernel": "k_sampler", "checksum": float(y.double().sum().item())}
    if a.cpu:
        import time
        from oracle import loader
        data = r.standard_normal(48000).astype(np.float32)
        out = np.zeros((F, 1), dtype=np.float32)
        t0 = time.perf_counter()
        loader.oracle().sampler(data, [(0, loader.EV_SAMPLE_MODE, 1, 0.0), (0, loader.EV_GATE_ON, 0, 0.0)], out)
        line["cpu_baseline"] = {"value": F / (time.perf_counter() - t0), "unit": "instance-samples/s", "cores": 1, "kind": "port",
                                "sample": f"1 instance x {F} frames, the oracle's Sample restatement"}
    print(json.dumps(line), flush=True)
    b.close()


if __name__ == "__main__":
    p = argparse.ArgumentParser()
    p.add_argument("unit", choices=["dattorro", "filterfx", "delayfx", "fxrack", "voices", "chorus", "sampler"])
    p.add_argument("--instances", type=int, default=8192)
    p.add_argument("--frames", type=int, default=48000)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--mode", choices=["exact", "fma"], default="exact", help="fx units: the arithmetic tier (olb_fx_set_mode)")
    a = p.parse_args()
    from ol_dsp_b200 import fx as _fx
    {"voices": bench_voices, "dattorro": bench_dattorro, "chorus": bench_chorus, "sampler": bench_sampler, "filterfx": lambda a: bench_fx(a, _fx.FILTER, "filterfx", 8),
     "delayfx": lambda a: bench_fx(a, _fx.DELAY, "delayfx", 16), "fxrack": lambda a: bench_fx(a, _fx.RACK, "fxrack", 16)}[a.unit](a)
