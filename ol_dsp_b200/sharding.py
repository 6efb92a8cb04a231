"""Multi-GPU host logic: units (channels / voices / instances) are independent, so rank r of
`world` owns one contiguous range and nothing crosses GPUs on the data path.  The single exchange step is the voice
bus: every rank holds a partial [frames][channels] fp32 bus in device memory and ONE in-place sum all-reduce produces
the global mixdown (Polyvoice.h:28-33 summed over all shards).  The reduce lives in libolbank.so (csrc/comm.cu: NCCL over
NVLink / NVSwitch, optionally on its own stream under the next block's voices); this file only carries the 128-byte id
from rank 0 to the other ranks over whatever process group the host already has, and marshals the calls."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check

ID_BYTES = 128


def shard_range(total, rank, world):
    """Contiguous, balanced: the first (total % world) ranks get one extra unit.  -> (first, count)"""
    assert 0 <= rank < world and total >= 0
    base, extra = divmod(total, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


class Comm:
    """One rank of the voice-bus communicator (olb_comm_*)."""

    def __init__(self, device, rank, world, id_bytes):
        self.L = _lib.lib()
        assert len(id_bytes) == ID_BYTES
        buf = (C.c_ubyte * ID_BYTES).from_buffer_copy(bytes(id_bytes))
        h = C.c_void_p()
        check(self.L.olb_comm_init(device, rank, world, C.cast(buf, C.c_void_p), C.byref(h)))
        self.h, self.rank, self.world = h, rank, world

    @staticmethod
    def unique_id():
        buf = (C.c_ubyte * ID_BYTES)()
        check(_lib.lib().olb_comm_unique_id(C.cast(buf, C.c_void_p)))
        return bytes(buf)

    @classmethod
    def from_torch_distributed(cls, device):
        """World = the default torch.distributed group (any backend): rank 0 makes the id, a broadcast carries it."""
        import torch
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()
        t = torch.zeros(ID_BYTES, dtype=torch.uint8)
        if rank == 0:
            t = torch.from_numpy(np.frombuffer(cls.unique_id(), dtype=np.uint8).copy())
        if dist.get_backend() == "nccl":
            t = t.cuda(device)
        dist.broadcast(t, 0)
        return cls(device, rank, world, t.cpu().numpy().tobytes())

    def close(self):
        if getattr(self, "h", None):
            self.L.olb_comm_destroy(self.h)
            self.h = None

    __del__ = close

    def AllReduceBus(self, bus_ptr, frames, channels, stream=0):
        check(self.L.olb_voice_bus_allreduce(self.h, C.c_void_p(bus_ptr), frames, channels, C.c_void_p(stream)))

    def AllReduceBusAsync(self, bus_ptr, frames, channels, producer_stream, slot):
        check(self.L.olb_voice_bus_allreduce_async(self.h, C.c_void_p(bus_ptr), frames, channels, C.c_void_p(producer_stream), slot))

    def Wait(self, slot, stream=0):
        check(self.L.olb_comm_wait(self.h, slot, C.c_void_p(stream)))

    def Sync(self):
        check(self.L.olb_comm_sync(self.h))
