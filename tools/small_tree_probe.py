#!/usr/bin/env python3
"""Byte-scan kernel time on small trees (the sizes move2kube sees in practice): 32 MB ... 1 GB, per library variant
(M2K_LIB selects).  Prints one line per size."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from move2kube_b200 import _abi, lib  # noqa: E402

ctx = lib.Context(0)
for mb, nf, nd in ((32, 2000, 120), (128, 4000, 250), (512, 8000, 500)):
    p = _abi.SynthParams(seed=42, n_files=nf, n_dirs=nd, max_depth=6, n_ignore_files=5, total_bytes=mb << 20)
    tree = ctx.synth_tree(p, host_bytes=False)
    dt = ctx.upload(tree)
    ctx.synth_fill_device(dt, p.seed)
    ms, tot = [], []
    for i in range(8):
        ctx.scan(dt)
        res = ctx.finalize(dt)
        ms.append(ctx.scan_kernel_ms())
        ph = ctx.phase_ms()
        tot.append(sum(ph[k] for k in ("names", "scan", "combine", "dockerfile", "finalize_kernels", "d2h")))
        ctx.result_free(res)
    ms, tot = sorted(ms[2:]), sorted(tot[2:])
    gb = tree.contents.content_bytes / 1e9
    print("%4d MB: scan %.3f ms = %.0f GB/s; device phases %.3f ms" % (mb, ms[len(ms) // 2], gb / ms[len(ms) // 2] * 1e3, tot[len(tot) // 2]), flush=True)
    ctx.devtree_free(dt)
    ctx.tree_free(tree)
