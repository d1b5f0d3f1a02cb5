"""Multi-GPU plumbing of the path: one process per GPU, one tiny exchange at the end.

Two exchanges exist.  PeerExchange (the default of bench.py) has no collective in the data path: the
ranks' kernels reduce straight into rank 0's arrays over NVLink (CUDA IPC peer memory, include/m2kdetect.h
m2k_exchange_*); torch.distributed only carries the 192-byte handle once, at set-up.  RankExchange gathers
with NCCL collectives and merges on rank 0; it is kept as the comparison and as the road for boxes where
peer memory between processes is not available.

The unit of work is the 64 KiB chunk: each rank scans a contiguous range of
chunks of near-equal bytes (lib.plan_shards / Context.upload_chunks); large files may be cut between
ranks, everything else goes whole to one rank; the directory table is replicated.  The only exchange
is the per-directory partials -- a content bit mask (OR) and the lowest main-file id (MIN) -- plus
two 32-byte piece summaries per rank, gathered with one all_gather (NCCL over NVLink on GPUs; NCCL has
no bitwise-OR reduction, so gather + merge kernels), and the per-file flag bytes (all_reduce).  The
rank that consumes the result merges and runs the cheap directory-level phases.
"""
import os
from bisect import bisect_left
from itertools import accumulate

from . import _abi, lib


def shard_ranges(file_len, world, weights=None):
    """Contiguous [begin, end) file-id ranges, one per rank, covering all files, near-equal in bytes --
    or, with `weights`, in proportion to them (e.g. the host-to-device rate each rank was measured at: on
    a box whose GPUs do not all reach the host equally fast, equal shards make everybody wait for the
    slowest link).  `file_len` is any sequence of file sizes."""
    n = len(file_len)
    if world <= 1:
        return [(0, n)]
    prefix = list(accumulate(file_len))
    total = prefix[-1] if prefix else 0
    w = [1.0] * world if not weights else [max(float(x), 1e-9) for x in weights]
    wsum = sum(w)
    cuts = [0]
    for r in range(1, world):
        target = total * sum(w[:r]) / wsum
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


def split_files(tree, cuts):
    """Files that the chunk cut points `cuts` (lib.plan_shards) divide between ranks."""
    total = cuts[-1]
    out = []
    for c in cuts[1:-1]:
        if 0 < c < total:
            f, k = lib.chunk_locate(tree, c)
            if k != 0 and f not in out:
                out.append(f)
    return out


def settle_undecided(ctx, tree, dt, files, fill_seed=None):
    """On the rank that finalises, after merge_pieces and before finalize: a file that was spread over
    several GPUs and still looks like a Dockerfile (M2K_FILE_UNDECIDED) gets one pass of its own -- the
    file is the unit again -- and its flags are written back.  Returns the files settled this way.
    `fill_seed`: the tree's bytes live only on the device (synthetic trees)."""
    settled = []
    for f in files:
        if not ctx.file_flag(dt, f) & _abi.FILE_UNDECIDED:
            continue
        one = ctx.upload(tree, f, f + 1)
        try:
            if fill_seed is not None:
                ctx.synth_fill_device(one, fill_seed)
            ctx.scan(one)
            ctx.settle_file(dt, f, ctx.file_flag(one, f) & 0xF)
        finally:
            ctx.devtree_free(one)
        settled.append(f)
    return settled


class RankExchange:
    """The one exchange step of a sharded pass, on torch.distributed (NCCL on GPUs): every rank
    contributes its per-directory partials, its two piece summaries (one all_gather) and its per-file
    flags (all_reduce: the ranges are disjoint, so the byte-wise sum is the union); rank 0 -- the one
    consumer of the result -- merges them into its device tree, ready for m2k_finalize.

    Nothing here waits on the host: the library's stream and torch's current stream (where the
    collectives run) are ordered against each other with events, so a rank queues its whole pass
    -- scan, export, collectives, merges -- and only m2k_finalize blocks."""

    def __init__(self, dist, torch, n_dirs, n_files, world, ctx, device="cuda"):
        self.dist, self.torch, self.nd, self.world = dist, torch, n_dirs, world
        self.lib_stream = torch.cuda.ExternalStream(ctx.stream())
        self.width = 2 * n_dirs + 16  # int32 words per rank: content mask, main-file id, 64 bytes of pieces
        self.mine = torch.empty((self.width,), dtype=torch.int32, device=device)
        self.all = torch.empty((world, self.width), dtype=torch.int32, device=device)
        self.allc = torch.empty((world, n_dirs), dtype=torch.int32, device=device)
        self.allm = torch.empty((world, n_dirs), dtype=torch.int32, device=device)
        self.allp = torch.empty((world, 16), dtype=torch.int32, device=device)
        self.flags = torch.empty((n_files,), dtype=torch.uint8, device=device)

    def gather(self, ctx, dt):
        nd, m = self.nd, self.mine
        cur = self.torch.cuda.current_stream()
        # the previous pass's collectives and merges (which a rank that does not finalise never waited
        # for) must be done with these buffers before the library writes them again
        cur.synchronize()
        ctx.export_async(dt, m[0:nd].data_ptr(), m[nd:2 * nd].data_ptr(), m[2 * nd:].data_ptr(), self.flags.data_ptr())
        cur.wait_stream(self.lib_stream)
        self.dist.all_gather_into_tensor(self.all.view(-1), m)
        self.dist.all_reduce(self.flags)

    def merge_on_root(self, ctx, dt, tree, split, fill_seed=None):
        """Rank 0 only, after gather()."""
        nd, a = self.nd, self.all
        self.allc.copy_(a[:, 0:nd])
        self.allm.copy_(a[:, nd:2 * nd])
        self.allp.copy_(a[:, 2 * nd:])
        self.lib_stream.wait_stream(self.torch.cuda.current_stream())
        ctx.import_async(dt, self.allc.data_ptr(), self.allm.data_ptr(), self.allp.data_ptr(), self.flags.data_ptr(), self.world)
        return settle_undecided(ctx, tree, dt, split, fill_seed) if split else []


class PeerExchange:
    """The exchange over peer memory (m2k_exchange_*): rank 0 opens it on its device tree and broadcasts
    the handle once; the other ranks attach.  After that a pass is: every rank m2k_scan (attached ranks
    reduce into rank 0's arrays from inside their kernels and raise a flag); rank 0 m2k_prepare,
    m2k_exchange_collect (a device-side wait + the stitching of cut files), m2k_finalize.  Nothing on
    the host of any rank waits for another rank."""

    def __init__(self, dist, torch, ctx, dt, rank, world, device="cuda"):
        self.rank, self.world = rank, world
        buf = torch.zeros((192,), dtype=torch.uint8, device=device)
        if rank == 0:
            h = ctx.exchange_open(dt, world)
            buf.copy_(torch.frombuffer(bytearray(h), dtype=torch.uint8))
        dist.broadcast(buf, 0)
        self.error = None
        if rank != 0:
            try:
                if os.environ.get("M2K_TEST_NO_IPC"):  # testing the fallback: behave as if the mapping had been refused
                    raise lib.M2KError(_abi.ERR_CUDA, "M2K_TEST_NO_IPC is set")
                ctx.exchange_attach(dt, rank, bytes(buf.cpu().numpy().tobytes()))
            except lib.M2KError as e:  # e.g. a sandbox that forbids CUDA IPC between processes
                self.error = str(e)
        # nobody scans before everybody is attached (pass numbers must stay in step), and everybody must
        # know whether everybody could attach
        ok = torch.tensor([0 if self.error else 1], dtype=torch.int32, device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        self.ok = bool(int(ok.item()))

    def after_scan(self, ctx, dt, tree=None, split=(), fill_seed=None, wait=True):
        """Rank 0: collect and settle (call between prepare and finalize).  Others: optionally block
        until this rank's part has been delivered."""
        if self.rank == 0:
            ctx.exchange_collect(dt)
            return settle_undecided(ctx, tree, dt, split, fill_seed) if split else []
        if wait == "previous":   # stay one pass behind: the next scan can be queued at once
            ctx.exchange_wait_previous(dt)
        elif wait:
            ctx.exchange_wait(dt)
        return []
