"""Worker of tests/test_gpu_multi.py: one process per GPU under torchrun.  Every rank scans its chunk
range of the same synthetic tree; rank 0 merges (partials, file flags, pieces), settles, finalises and
compares every array with a single-GPU pass over the whole tree."""
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
    for split_min in (1 << 20, 1 << 40):      # large files cut between ranks / files kept whole
        cuts = lib.plan_shards(tree, world, split_min)
        split = multi.split_files(tree, cuts)
        dt = ctx.upload_chunks(tree, cuts[rank], cuts[rank + 1])
        xch = multi.RankExchange(dist, torch, nd, nf, world)
        for _ in range(2):                    # twice: the second pass reuses every buffer
            ctx.scan(dt)
            if rank == 0:
                ctx.prepare(dt)
            xch.gather(ctx, dt)
            if rank == 0:
                xch.merge_on_root(ctx, dt, tree, split)
                res = ctx.finalize(dt)
                whole = ctx.detect(tree)
                g, w = res.contents, whole.contents
                same = g.file_flags[0:nf] == w.file_flags[0:nf] and g.n_services == w.n_services
                for name in ("dir_state", "dir_opts", "dir_main_file", "dir_subtree", "dir_direct", "dir_ignore"):
                    same = same and getattr(g, name)[0:nd] == getattr(w, name)[0:nd]
                same = same and g.service_dirs[0:g.n_services] == w.service_dirs[0:w.n_services]
                print("split_min=%d cut_files=%d services=%d same=%s" % (split_min, len(split), g.n_services, same), flush=True)
                ok = ok and same and (split_min > (1 << 30) or len(split) > 0)
                ctx.result_free(res)
                ctx.result_free(whole)
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
