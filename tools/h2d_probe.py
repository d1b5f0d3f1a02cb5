#!/usr/bin/env python3
"""What limits the end-to-end number when several GPUs copy from the host at once?

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/h2d_probe.py [--gb 4]

Every rank page-locks `--gb` GB of host memory and copies it to its GPU with plain cudaMemcpyAsync (one copy per
32 MiB block, nothing else running), all ranks at the same time; then again with a single rank copying alone.
Two placements of the host side are compared:
  unbound : threads and pages wherever the kernel puts them (what bench.py's e2e leg does)
  bound   : before allocating, each rank pins itself to its own slice of the host's cores
            (sched_setaffinity) and first-touches its buffer from there
If `bound` recovers the single-rank bandwidth, placement was the limit and the library should place its pinned
blocks; if both stay low, the host's memory / PCIe root complex is the ceiling and the e2e curve reports the box.
Rank 0 prints one JSON line.
"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist


def copy_rate(buf, dev, reps, block=32 << 20):
    n = buf.numel()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for off in range(0, n, block):
            dev[off:off + block].copy_(buf[off:off + block], non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    return n * reps / 1e9 / (e0.elapsed_time(e1) / 1e3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=4.0)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = int(args.gb * (1 << 30)) >> 25 << 25
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    all_cores = sorted(os.sched_getaffinity(0))
    out = {"ranks": world, "gb_per_rank": nbytes / 1e9, "host_cores": len(all_cores), "placements": {}}
    for placement in ("unbound", "bound"):
        if placement == "bound":
            per = max(1, len(all_cores) // world)
            os.sched_setaffinity(0, set(all_cores[rank * per:(rank + 1) * per]) or set(all_cores))
        t0 = time.perf_counter()
        buf = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        buf.fill_(rank + 1)  # first touch from this rank's cores
        t_pin = time.perf_counter() - t0
        copy_rate(buf, dev, 1)  # warm-up
        if world > 1:
            dist.barrier()
        together = copy_rate(buf, dev, args.reps)
        rates = torch.tensor([together], device="cuda", dtype=torch.float64)
        alone = torch.zeros(1, device="cuda", dtype=torch.float64)
        if world > 1:
            gathered = [torch.zeros_like(rates) for _ in range(world)]
            dist.all_gather(gathered, rates)
            dist.barrier()
            if rank == 0:  # one rank alone, the others idle
                alone[0] = copy_rate(buf, dev, args.reps)
            dist.barrier()
            per_rank = [float(g.item()) for g in gathered]
        else:
            per_rank = [together]
            alone[0] = together
        out["placements"][placement] = {"per_rank_GBps_all_copying": per_rank, "sum_GBps": sum(per_rank),
                                        "rank0_alone_GBps": float(alone.item()), "pin_and_touch_s": t_pin}
        del buf
        os.sched_setaffinity(0, set(all_cores))
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
