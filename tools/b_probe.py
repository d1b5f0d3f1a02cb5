#!/usr/bin/env python3
"""Config B (10k files, ~1 GB) pass by pass: device phase times, for profiling the small-tree path
(`ncu -k regex:k_df ... python tools/b_probe.py 3`)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from move2kube_b200 import _abi, lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 12
ctx = lib.Context(0)
p = _abi.SynthParams(seed=1, n_files=10_000, n_dirs=600, max_depth=6, n_ignore_files=20, total_bytes=1_000_000_000)
tree = ctx.synth_tree(p, host_bytes=False)
dt = ctx.upload(tree)
ctx.synth_fill_device(dt, p.seed)
rows = []
for i in range(n):
    ctx.scan(dt)
    res = ctx.finalize(dt)
    rows.append(dict(ctx.phase_ms()))
    ctx.result_free(res)
keys = [k for k in rows[-1]]
rows = rows[2:] or rows
med = {k: sorted(r[k] for r in rows)[len(rows) // 2] for k in keys}
print("config B phases (median ms): " + ", ".join("%s %.3f" % (k, med[k]) for k in keys), flush=True)
ctx.devtree_free(dt)
ctx.tree_free(tree)
