"""Multi-GPU plumbing of the path: one process per GPU, files sharded, one tiny gather at the end.

Files are independent units: each rank scans a contiguous range of file ids
chosen so that the ranges carry near-equal bytes; the directory table is replicated.  The only
exchange is the per-directory partials -- a content bit mask (OR) and the lowest main-file id (MIN),
2 x n_dirs uint32 per rank -- gathered with one all_gather (NCCL over NVLink on GPUs; NCCL has no
bitwise-OR reduction, so gather + a merge kernel) and folded on every rank, after which the cheap
directory-level phases run replicated.
"""
from bisect import bisect_left
from itertools import accumulate


def shard_ranges(file_len, world):
    """Contiguous [begin, end) file-id ranges, one per rank, covering all files, near-equal in bytes.
    `file_len` is any sequence of file sizes."""
    n = len(file_len)
    if world <= 1:
        return [(0, n)]
    prefix = list(accumulate(file_len))
    total = prefix[-1] if prefix else 0
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        c = bisect_left(prefix, target) + 1 if total else n * r // world
        cuts.append(min(max(c, cuts[-1]), n))
    cuts.append(n)
    return [(cuts[i], cuts[i + 1]) for i in range(world)]


def gather_partials(dist, mine, out=None):
    """all_gather of this rank's [2, n_dirs] int32 partials into [world, 2, n_dirs]."""
    import torch
    world = dist.get_world_size()
    if out is None:
        out = torch.empty((world,) + tuple(mine.shape), dtype=mine.dtype, device=mine.device)
    dist.all_gather_into_tensor(out.view(-1), mine.contiguous().view(-1))
    return out
