"""Image-batch sharding across replicas (one process per GPU). The PyraPose hot path has no data-path
collective: images (and anchor-target work items) are independent, so each rank takes a contiguous
slice and results are concatenated on the host. torch.distributed is used only for
the barrier, the max-over-ranks reduction of timings and the optional gather of the small results."""
import numpy as np


def shard_range(total, rank, world):
    """Contiguous [begin, end) slice of `total` items owned by `rank`; sizes differ by at most one and
    lower ranks take the remainder (the same rule numpy.array_split uses)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world: %r/%r" % (rank, world))
    if total < 0:
        raise ValueError("total must be >= 0")
    base, rem = divmod(total, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def max_over_ranks(value, device=None):
    """max of a python float over all ranks (no-op without an initialised process group)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_arrays(arrays, dst=0):
    """Concatenates each rank's list of numpy arrays along axis 0 on rank `dst` (rank order = image order).
    Returns the list of concatenated arrays on `dst`, None elsewhere."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [np.asarray(a) for a in arrays]
    world, rank = dist.get_world_size(), dist.get_rank()
    bucket = [None] * world if rank == dst else None
    dist.gather_object([np.asarray(a) for a in arrays], bucket, dst=dst)
    if rank != dst:
        return None
    return [np.concatenate([bucket[r][i] for r in range(world)], axis=0) for i in range(len(arrays))]
