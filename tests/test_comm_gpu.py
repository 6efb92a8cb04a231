"""GPU tests of the cross-GPU voice bus: the NCCL all-reduce behind the C ABI
(olb_comm_* / olb_voice_bus_allreduce*), with the voices of a Polyvoice sharded over the ranks.
world = 1 runs on any box; world = 2 needs two GPUs (`gpurun --gpus 2`) and is skipped otherwise."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-5


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _notes(voices):
    return [36 + v % 61 for v in range(voices)]


def _rank(rank, world, port, voices, frames, q):
    """One process per GPU: its shard of the voices -> partial bus -> olb_voice_bus_allreduce -> every rank holds the mix."""
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist

    from ol_dsp_b200 import Comm, VoiceBank, shard_range
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)   # plumbing only: carries the 128-byte id
    comm = Comm.from_torch_distributed(rank)
    first, count = shard_range(voices, rank, world)
    b = VoiceBank(count, device=rank)
    b.NoteOn(np.asarray(_notes(voices)[first:first + count], dtype=np.uint8))
    st = torch.cuda.current_stream().cuda_stream
    half = frames // 2
    # first half: the in-stream form; second half: the overlapped form into a second bus buffer
    bus = torch.zeros((frames, 2), dtype=torch.float32, device="cuda")
    b.ProcessDevice(half, bus_ptr=bus.data_ptr(), bus_channels=2, bus_gain=1.0 / voices, stream=st)
    comm.AllReduceBus(bus.data_ptr(), half, 2, st)
    b.ProcessDevice(frames - half, bus_ptr=bus[half:].data_ptr(), bus_channels=2, bus_gain=1.0 / voices, stream=st)
    comm.AllReduceBusAsync(bus[half:].data_ptr(), frames - half, 2, st, 1)
    comm.Wait(1, st)
    torch.cuda.synchronize()
    q.put((rank, bus.cpu().numpy()))
    dist.barrier()
    comm.close()
    b.close()
    dist.destroy_process_group()


def test_comm_world_one_is_the_identity_and_orders_streams(orc):
    """A one-rank communicator: the reduce leaves the bus as it is; the asynchronous form runs on the communicator's stream
    after the producer and olb_comm_wait orders the consumer behind it."""
    from ol_dsp_b200 import Comm, VoiceBank
    comm = Comm(0, 0, 1, Comm.unique_id())
    V, F = 64, 3000
    b = VoiceBank(V)
    b.NoteOn(np.asarray(_notes(V), dtype=np.uint8))
    s = torch.cuda.Stream()
    bus = torch.zeros((F, 2), dtype=torch.float32, device="cuda")
    ref = torch.zeros((F, 2), dtype=torch.float32, device="cuda")
    b.ProcessDevice(F, bus_ptr=bus.data_ptr(), bus_channels=2, bus_gain=1.0 / V, stream=s.cuda_stream)
    comm.AllReduceBusAsync(bus.data_ptr(), F, 2, s.cuda_stream, 0)
    comm.Wait(0, torch.cuda.current_stream().cuda_stream)
    got = bus.clone()  # on the current stream, behind the reduce
    torch.cuda.synchronize()
    b.Init(48000.0)
    b.NoteOn(np.asarray(_notes(V), dtype=np.uint8))
    b.ProcessDevice(F, bus_ptr=ref.data_ptr(), bus_channels=2, bus_gain=1.0 / V)
    torch.cuda.synchronize()
    assert torch.equal(got, ref) and float(ref.abs().max()) > 0
    from oracle import loader
    ev = [(0, loader.EV_NOTE_ON, n, 100, v) for v, n in enumerate(_notes(V))]
    want = orc.polyvoice(V, F, 48000.0, None, ev)
    assert np.abs(got.cpu().numpy()[:, 0] - want / V).max() <= TOL  # bus = mean of the voices: full scale stays ~1
    comm.close()
    b.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_sharded_polyvoice_bus_over_two_gpus_matches_the_oracle(orc):
    """Voices [0, V/2) on GPU 0, [V/2, V) on GPU 1; after the all-reduce BOTH ranks hold the mix of all V voices, equal to
    the reference's Polyvoice sum (oracle) within the summation-order tolerance."""
    import torch.multiprocessing as mp

    from oracle import loader
    voices, frames, world = 96, 4000, 2
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_rank, args=(r, world, port, voices, frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get() for _ in range(world))
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    ev = [(0, loader.EV_NOTE_ON, n, 100, v) for v, n in enumerate(_notes(voices))]
    want = orc.polyvoice(voices, frames, 48000.0, None, ev)
    assert np.array_equal(got[0], got[1])  # one deterministic reduce: every rank holds the same bits
    for ch in (0, 1):
        assert np.abs(got[0][:, ch] - want / voices).max() <= TOL
