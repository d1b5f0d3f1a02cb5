"""CPU tests of the N>1 plumbing: shard planning, and a world_size-2 gloo run of the gather with the
partials of each shard computed by the oracle (the GPU kernels are covered by -m gpu)."""
import os
import random
import subprocess
import sys
import textwrap

from move2kube_b200 import multi

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_cover_everything_and_balance_bytes():
    rng = random.Random(1)
    for world in (1, 2, 3, 4, 8):
        lens = [int(rng.lognormvariate(10, 1.5)) for _ in range(5000)]
        r = multi.shard_ranges(lens, world)
        assert len(r) == world and r[0][0] == 0 and r[-1][1] == len(lens)
        assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
        total = sum(lens)
        for a, b in r:
            assert abs(sum(lens[a:b]) - total / world) <= max(lens) + 1
    # skew: a few huge files and many tiny ones (config D); files are the unit, so a rank may own one huge file
    lens = [4 << 30] * 8 + [1000] * 100000
    r = multi.shard_ranges(lens, 8)
    assert sorted(b - a for a, b in r)[0] >= 1 and r[-1][1] == len(lens)
    assert multi.shard_ranges([], 4) == [(0, 0)] * 4
    assert multi.shard_ranges([0, 0, 0], 2)[-1][1] == 3


WORKER = textwrap.dedent("""
    import os, sys
    sys.path[:0] = [%(repo)r, os.path.join(%(repo)r, "tests")]
    import torch, torch.distributed as dist
    import oracle_lib
    from move2kube_b200 import _abi, lib, multi
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%(port)d", rank=int(sys.argv[1]), world_size=2)
    rank = dist.get_rank()
    p = _abi.SynthParams(seed=9, n_files=1200, n_dirs=90, max_depth=5, n_ignore_files=4, total_bytes=6 << 20)
    tree = lib.synth_tree_host(p, host_bytes=True)
    t = tree.contents
    nd, nf = t.n_dirs, t.n_files
    a, b = multi.shard_ranges(t.file_len[0:nf], 2)[rank]
    flags = (oracle_lib.C.c_uint8 * max(1, b - a))()
    assert oracle_lib.lib().m2ko_scan_files(tree, a, b, flags, 2) == 0
    mine = torch.zeros((2, nd), dtype=torch.int32)
    mine[1].fill_(-1)  # M2K_NONE as int32
    names = bytes(bytearray(t.names[0:t.n_name_bytes]))
    for f in range(a, b):
        e = t.file_entry[f]
        d = t.entry_dir[e]
        nm = names[t.entry_name_off[e]:t.entry_name_off[e + 1]]
        if flags[f - a] & _abi.FILE_HAS_MAIN and nm.endswith(b"py"):
            mine[0, d] |= 1 << _abi.C_MAIN_PY
            cur = mine[1, d].item() & 0xFFFFFFFF
            mine[1, d] = min(cur, f) if cur != 0xFFFFFFFF else f
        if flags[f - a] & _abi.FILE_DOCKERFILE:
            mine[0, d] |= 1 << _abi.C_DOCKERFILE
    allp = multi.gather_partials(dist, mine)
    content = allp[:, 0, :]
    merged_content = content[0] | content[1]
    mains = allp[:, 1, :].to(torch.int64) & 0xFFFFFFFF
    merged_main = mains.min(dim=0).values
    # the whole-tree answer from the oracle: direct (not subtree) values
    res = oracle_lib.detect(tree, n_threads=2)
    r = res.contents
    want_c = [0] * nd
    want_m = [0xFFFFFFFF] * nd
    for f in range(nf):
        e = t.file_entry[f]
        d = t.entry_dir[e]
        nm = names[t.entry_name_off[e]:t.entry_name_off[e + 1]]
        if r.file_flags[f] & _abi.FILE_HAS_MAIN and nm.endswith(b"py"):
            want_c[d] |= 1 << _abi.C_MAIN_PY
            want_m[d] = min(want_m[d], f)
        if r.file_flags[f] & _abi.FILE_DOCKERFILE:
            want_c[d] |= 1 << _abi.C_DOCKERFILE
    assert merged_content.tolist() == want_c
    assert merged_main.tolist() == want_m
    assert sum(want_c) > 0
    dist.barrier()
    dist.destroy_process_group()
    print("rank", rank, "ok")
""")


def test_two_rank_gather_on_gloo(tmp_path, oracle):
    port = 29000 + os.getpid() % 2000
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"repo": REPO, "port": port})
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, p in enumerate(procs):
        assert p.returncode == 0, outs[r][-2000:]
        assert "rank %d ok" % r in outs[r]


def test_weighted_shard_ranges_follow_the_weights():
    """bench.py's e2e leg deals files out in proportion to each rank's measured host-to-device rate."""
    lens = [1000] * 10000
    r = multi.shard_ranges(lens, 4, [23, 23, 34.5, 34.5])
    sizes = [b - a for a, b in r]
    assert r[0][0] == 0 and r[-1][1] == len(lens) and all(r[i][1] == r[i + 1][0] for i in range(3))
    assert sizes[0] == sizes[1] and sizes[2] == sizes[3] and abs(sizes[2] / sizes[0] - 1.5) < 0.01
    assert multi.shard_ranges(lens, 4, [1, 1, 1, 1]) == multi.shard_ranges(lens, 4)
    assert multi.shard_ranges(lens, 2, [0, 1])[0] == (0, 1) or multi.shard_ranges(lens, 2, [0, 1])[0][1] <= 1


def test_bench_snapshot_comparison_counts_differences():
    import numpy as np
    sys.path.insert(0, REPO)
    import bench
    a = {k: np.arange(10, dtype=np.uint32) for k in bench.RESULT_ARRAYS}
    b = {k: v.copy() for k, v in a.items()}
    assert bench.compare_snapshots(a, b) == {}
    b["dir_opts"][3] = 99
    b["service_dirs"] = b["service_dirs"][:8]
    d = bench.compare_snapshots(a, b)
    assert d == {"dir_opts": 1, "service_dirs": 2}
