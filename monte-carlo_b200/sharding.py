"""Multi-GPU sharding of the path: every unit (hand, combination index, board index, Philox
trial) is independent, so ranks take contiguous index ranges and exchange nothing but one all-reduce of
the int64 counts at the end.  One process per GPU; torch.distributed is only the plumbing."""


def shard_range(total, rank, world):
    """contiguous slice [lo, hi) of [0, total) for `rank` of `world`; slices tile the range exactly"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return total * rank // world, total * (rank + 1) // world


def allreduce_counts(counts, dist=None):
    """in-place SUM all-reduce of an int64 tensor of counts (NCCL over NVLink on GPUs, gloo in CPU tests)"""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    return counts
