"""Worker of tests/test_gpu_multi.py: one process per GPU under torchrun.  Every rank scans its chunk
range of the same synthetic tree; rank 0 merges (partials, file flags, pieces), settles, finalises and
compares every array with a single-GPU pass over the whole tree.  Both exchanges are run: over peer
memory (m2k_exchange_*, CUDA IPC) and over NCCL collectives (M2K_EXCHANGE=peer,nccl selects)."""
import os
import sys

import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from move2kube_b200 import _abi, lib, multi  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    import datetime
    dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=120))
    ctx = lib.Context(local)
    p = _abi.SynthParams(seed=33, n_files=3000, n_dirs=200, max_depth=6, n_ignore_files=6, total_bytes=160 << 20,
                         n_huge=3, huge_bytes=24 << 20, n_large=3, large_bytes=5 << 20)
    tree = ctx.synth_tree(p, host_bytes=True)
    t = tree.contents
    nd, nf = t.n_dirs, t.n_files
    ok = True
    lens = t.file_len[0:nf]
    biggest = max(range(nf), key=lambda f: lens[f])
    before = sum((n + 65535) // 65536 for n in lens[0:biggest])
    total = lib.chunk_count(tree)
    kinds = [k for k in os.environ.get("M2K_EXCHANGE", "peer,nccl").split(",") if k]
    for split_min in (1 << 20, 1 << 40, 0):   # the planner with / without cutting large files; cuts forced through the largest file
        if split_min:
            cuts = lib.plan_shards(tree, world, split_min)
        else:
            nb = (lens[biggest] + 65535) // 65536
            inner = sorted(before + nb * (r + 1) // (world + 1) for r in range(world - 1))
            cuts = [0] + inner + [total]
        split = multi.split_files(tree, cuts)
        for kind in kinds:
            dt = ctx.upload_chunks(tree, cuts[rank], cuts[rank + 1])
            if kind == "peer":
                xch = multi.PeerExchange(dist, torch, ctx, dt, rank, world)
            else:
                xch = multi.RankExchange(dist, torch, nd, nf, world, ctx)
            for _ in range(2):                    # twice: the second pass reuses every buffer
                ctx.scan(dt)
                if rank == 0:
                    ctx.prepare(dt)
                if kind == "peer":
                    xch.after_scan(ctx, dt, tree, split)
                else:
                    xch.gather(ctx, dt)
                if rank == 0:
                    if kind != "peer":
                        xch.merge_on_root(ctx, dt, tree, split)
                    res = ctx.finalize(dt)
                    whole = ctx.detect(tree)
                    g, w = res.contents, whole.contents
                    diffs = []
                    for name, n in (("file_flags", nf), ("dir_state", nd), ("dir_opts", nd), ("dir_main_file", nd), ("dir_subtree", nd),
                                    ("dir_direct", nd), ("dir_ignore", nd), ("service_dirs", min(g.n_services, w.n_services))):
                        a, b = getattr(g, name)[0:n], getattr(w, name)[0:n]
                        if a != b:
                            i = next(k for k in range(n) if a[k] != b[k])
                            diffs.append("%s[%d]: %r != %r (%d differ)" % (name, i, a[i], b[i], sum(1 for k in range(n) if a[k] != b[k])))
                    same = not diffs and g.n_services == w.n_services
                    if diffs:
                        print("DIFF", "; ".join(diffs), flush=True)
                    print("exchange=%s split_min=%d cut_files=%d services=%d same=%s" % (kind, split_min, len(split), g.n_services, same), flush=True)
                    ok = ok and same and (split_min != 0 or len(split) > 0 or world == 1)
                    ctx.result_free(res)
                    ctx.result_free(whole)
            dist.barrier()  # the root's arrays stay mapped on the other ranks until everybody is through
            ctx.devtree_free(dt)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.barrier()
    dist.destroy_process_group()
    ctx.tree_free(tree)
    ctx.close()
    if rank == 0:
        print("MGPU_OK" if ok else "MGPU_MISMATCH", flush=True)
    return 0 if int(flag.item()) else 1


if __name__ == "__main__":
    sys.exit(main())
