"""Multi-GPU: images are independent, so a batch is cut into contiguous image-index ranges, one per
rank (one process per GPU), with NO collective on the data path.  torch.distributed
carries only the bench's barrier and the max-over-ranks of the timed region."""


def shard_range(total, rank, world):
    """[lo, hi) of `total` images for `rank` of `world`: contiguous, disjoint, sizes differ by <= 1."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world %d" % (rank, world))
    q, r = divmod(total, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def weak_shard(per_gpu, rank, base=1000):
    """Weak scaling (bench.py): every rank decodes `per_gpu` images of its own; returns its seed range."""
    lo = base + rank * per_gpu
    return lo, lo + per_gpu


def max_over_ranks(x, device=None):
    """Time-like scalars are reported as the maximum over ranks (the job ends when the slowest rank does)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
