"""The N > 1 host logic on CPU (gloo, world_size 2): contiguous sharding of the units, the 128-byte communicator id
carried from rank 0 to the others over the host's process group, and the arithmetic of the one exchange step of the hot
path, the voice-bus sum.  Each rank computes the partial bus of ITS voices with the oracle (the
CUDA bank and its NCCL reduce need GPUs: tests/test_voice_gpu.py), the partial buses are summed, and rank 0 compares
with the oracle's single-process Polyvoice."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ol_dsp_b200.sharding import ID_BYTES, shard_range


def test_shard_range_covers_everything_once():
    for total in (0, 1, 7, 64, 65536, 65537):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            for (f0, c0), (f1, _) in zip(spans[:-1], spans[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, voices, frames, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import loader
    orc = loader.oracle()
    first, count = shard_range(voices, rank, world)
    bus = np.zeros(frames, dtype=np.float32)
    for v in range(first, first + count):  # every voice of the shard plays its own note from frame 0
        bus += orc.voice(frames, 48000.0, None, [(0, loader.EV_NOTE_ON, 36 + v % 61, 100)])
    # the id broadcast of Comm.from_torch_distributed (the bytes only: olb_comm_init itself needs a GPU)
    ident = torch.arange(ID_BYTES, dtype=torch.uint8) if rank == 0 else torch.zeros(ID_BYTES, dtype=torch.uint8)
    dist.broadcast(ident, 0)
    assert ident.tolist() == list(range(ID_BYTES))
    t = torch.from_numpy(bus)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)  # what olb_voice_bus_allreduce computes on the GPUs (fp32 sum over ranks)
    if rank == 0:
        q.put(t.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_voice_bus_allreduce_over_two_ranks_matches_single_process(orc):
    from oracle import loader
    voices, frames, world = 6, 2000, 2
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, voices, frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get()
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    ev = [(0, loader.EV_NOTE_ON, 36 + v % 61, 100, v) for v in range(voices)]
    want = orc.polyvoice(voices, frames, 48000.0, None, ev)
    assert np.abs(got - want).max() <= 1e-5  # same voices, summation order differs across shards
