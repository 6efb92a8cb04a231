"""Multi-GPU host logic: units (channels / voices / instances) are independent, so rank r of
`world` owns one contiguous range and nothing crosses GPUs on the data path.  The single exchange step is the voice
bus: every rank holds a partial [frames][channels] fp32 bus and ONE sum all-reduce (NCCL over NVLink on the GPU box,
gloo in the CPU tests) produces the global mixdown (Polyvoice.h:28-33 summed over all shards)."""


def shard_range(total, rank, world):
    """Contiguous, balanced: the first (total % world) ranks get one extra unit.  -> (first, count)"""
    assert 0 <= rank < world and total >= 0
    base, extra = divmod(total, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def reduce_bus(bus, group=None):
    """In-place sum of the partial buses over the ranks of `group` (default: world).  A no-op without a process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(bus, op=dist.ReduceOp.SUM, group=group)
    return bus
