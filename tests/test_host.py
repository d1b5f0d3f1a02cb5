"""CPU tests of the boundary's host logic: the C ABI library loads and exports what the header
declares, the product packer agrees with the oracle's independent walk, synthetic trees are
well formed and deterministic, path helpers follow Go.  No compute entry point is called."""
import ctypes as C
import os
import re
import subprocess

import pytest

import golden_check as gc
import treespec
from move2kube_b200 import _abi, lib

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDENS = gc.load_goldens()


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(REPO, "include", "m2kdetect.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(m2k_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    L = lib.load()
    missing = [n for n in sorted(declared) if not hasattr(L, n)]
    assert not missing, "declared in include/m2kdetect.h but not exported: %s" % missing
    assert set(lib.EXPORTS) == declared
    assert L.m2k_abi_version() == 1


def test_header_compiles_as_c(tmp_path):
    """cgo compiles include/m2kdetect.h as C, not C++: it must be valid C11 on its own."""
    import subprocess
    src = tmp_path / "hdr.c"
    src.write_text('#include "m2kdetect.h"\nint main(void) { m2k_tree t; (void)t; return M2K_ABI_VERSION - 1; }\n')
    subprocess.check_call(["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-pedantic", "-fsyntax-only",
                           "-I", os.path.join(REPO, "include"), str(src)])


def test_no_gpu_means_a_loud_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(lib.M2KError) as ei:
        lib.Context(0)
    assert ei.value.code == _abi.ERR_CUDA
    assert "no CPU path" in ei.value.message


def _tree_arrays(tp):
    t = tp.contents
    nd, ne, nf = t.n_dirs, t.n_entries, t.n_files
    d = {
        "n": (nd, ne, nf, t.max_depth, t.content_bytes),
        "dir_parent": t.dir_parent[0:nd], "dir_end": t.dir_end[0:nd], "dir_entry": t.dir_entry[0:nd],
        "dir_first_entry": t.dir_first_entry[0:nd + 1], "dir_depth": t.dir_depth[0:nd], "dir_flags": t.dir_flags[0:nd],
        "entry_dir": t.entry_dir[0:ne], "entry_name_off": t.entry_name_off[0:ne + 1], "entry_type": t.entry_type[0:ne],
        "entry_child": t.entry_child[0:ne], "names": bytes(bytearray(t.names[0:t.n_name_bytes])),
        "file_entry": t.file_entry[0:nf], "file_off": t.file_off[0:nf], "file_len": t.file_len[0:nf],
    }
    if t.bytes:
        d["bytes"] = bytes(bytearray(t.bytes[0:t.n_bytes]))
    return d


@pytest.mark.parametrize("spec", [g for g in GOLDENS if not g.get("nonexistent")], ids=lambda g: g["name"])
def test_packer_matches_the_oracle_walk(oracle, tmp_path, spec):
    root = os.fsencode(treespec.materialize(spec, str(tmp_path / spec["name"])))
    rc, otree = oracle.pack_fs(root)
    assert rc == 0
    ptree = lib.pack_tree_host(root)
    try:
        a, b = _tree_arrays(otree), _tree_arrays(ptree)
        for k in a:
            assert a[k] == b[k], k
        assert ptree.contents.root_path == root
    finally:
        lib.tree_free_host(ptree)
        oracle.lib().m2ko_tree_free(otree)


def test_packer_errors(tmp_path):
    with pytest.raises(lib.M2KError) as ei:
        lib.pack_tree_host(os.fsencode(str(tmp_path / "nope")))
    assert ei.value.code == _abi.ERR_IO
    # filepath.Walk Lstats its root: a file or a symlink (even to a directory) has no packed form
    (tmp_path / "real").mkdir()
    os.symlink(str(tmp_path / "real"), str(tmp_path / "link"))
    (tmp_path / "file").write_text("x")
    for name in ("link", "file"):
        with pytest.raises(lib.M2KError) as ei:
            lib.pack_tree_host(os.fsencode(str(tmp_path / name)))
        assert ei.value.code == _abi.ERR_NOTDIR
    d = tmp_path / "nl"
    d.mkdir()
    (d / "a\nb.go").write_text("x")
    with pytest.raises(lib.M2KError) as ei:
        lib.pack_tree_host(os.fsencode(str(d)))
    assert ei.value.code == _abi.ERR_UNSUPPORTED


def small_synth(seed=7, n_files=400, n_dirs=50, total=3 << 20, **kw):
    return _abi.SynthParams(seed=seed, n_files=n_files, n_dirs=n_dirs, max_depth=5, n_ignore_files=4,
                            total_bytes=total, n_huge=kw.get("n_huge", 0), huge_bytes=kw.get("huge_bytes", 0),
                            tiny_min=kw.get("tiny_min", 0), tiny_max=kw.get("tiny_max", 0))


def test_synthetic_tree_is_well_formed_and_deterministic(oracle):
    t1 = lib.synth_tree_host(small_synth())
    t2 = lib.synth_tree_host(small_synth())
    try:
        a, b = _tree_arrays(t1), _tree_arrays(t2)
        assert a == b
        t = t1.contents
        nd, ne, nf = t.n_dirs, t.n_entries, t.n_files
        assert nd == 50 and nf == 400 and ne == nd - 1 + nf
        view = gc.TreeView(t1)
        # pre-order ids with children in byte order; entries sorted by (dir, name)
        for d in range(1, nd):
            p = t.dir_parent[d]
            assert p < d and t.dir_depth[d] == t.dir_depth[p] + 1 and p <= d < t.dir_end[p]
        for e in range(1, ne):
            assert (t.entry_dir[e - 1], view.entry_name[e - 1]) < (t.entry_dir[e], view.entry_name[e])
        for f in range(nf):
            assert t.file_off[f] % 16 == 0
            if f:
                assert t.file_off[f] == t.file_off[f - 1] + (t.file_len[f - 1] + 15) // 16 * 16
        assert abs(t.content_bytes - (3 << 20)) < (3 << 20) * 0.05
        assert sum(1 for n in view.entry_name if n == b".m2kignore") >= 1
    finally:
        lib.tree_free_host(t1)
        lib.tree_free_host(t2)


def test_synthetic_tree_exercises_every_signature(oracle):
    """The oracle accepts the layout and the tree contains every case the kernels must tell apart."""
    tp = lib.synth_tree_host(small_synth(seed=7, n_files=4000, n_dirs=200, total=16 << 20))
    try:
        nf, nd = tp.contents.n_files, tp.contents.n_dirs
        res = oracle.detect(tp, n_threads=4)
        r = res.contents
        flags = r.file_flags[0:nf]
        states = set(r.dir_state[0:nd])
        assert r.n_services > 3
        assert {_abi.DIR_NOT_VISITED, _abi.DIR_IGNORED_DESCEND, _abi.DIR_NOMATCH_DESCEND, _abi.DIR_SERVICE} <= states
        assert any(x & _abi.FILE_DOCKERFILE for x in flags)
        assert any(x & _abi.FILE_HAS_MAIN for x in flags)
        assert any((x & _abi.FILE_FROM_FIRST) and not (x & _abi.FILE_DOCKERFILE) for x in flags)
        assert any(m != _abi.NONE for m in r.dir_main_file[0:nd])
        oracle.lib().m2ko_result_free(res)
    finally:
        lib.tree_free_host(tp)


def test_synth_fill_host_is_random_access():
    tp = lib.synth_tree_host(small_synth())
    try:
        t = tp.contents
        whole = bytes(bytearray(t.bytes[0:t.n_bytes]))
        f0, f1 = 100, 180
        lo = t.file_off[f0]
        hi = t.file_off[f1 - 1] + (t.file_len[f1 - 1] + 15) // 16 * 16
        buf = (C.c_uint8 * (hi - lo))()
        rc = lib.load().m2k_synth_fill_host(tp, 7, f0, f1, buf, 3)
        assert rc == 0
        assert bytes(buf) == whole[lo:hi]
        assert lib.load().m2k_synth_fill_host(tp, 8, f0, f1, buf, 3) == _abi.ERR_ARG
    finally:
        lib.tree_free_host(tp)


def test_path_helpers_follow_go(oracle):
    import gopath
    for p in [b"", b"/", b"a//b/../c/", b"../x/..", b"/a/../../b", b"./", b"a/b/../../.."]:
        assert lib.path_clean(p) == gopath.clean(p) == oracle.clean(p)
    for a, b in [(b"/R/K", "/r/K".encode()), (b"ss", "ß".encode()), (b"a\xffb", b"A\xfeB"), (b"I", "ı".encode()),
                 ("ſ".encode(), b"S"), (b"abc", b"ab")]:
        assert lib.equal_fold(a, b) == gopath.equal_fold(a, b) == oracle.equal_fold(a, b)


def _walk_order(a):
    """filepath.Walk order of all entry ids, straight from the packed arrays (the definition: a
    directory's entries in name order, descending into each real directory right after naming it)."""
    out = []

    def visit(d):
        for e in range(a["dir_first_entry"][d], a["dir_first_entry"][d + 1]):
            out.append(e)
            if (a["entry_type"][e] & 0xF) == _abi.T_DIR and a["entry_child"][e] != _abi.NONE:
                visit(a["entry_child"][e])
    visit(0)
    return out


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_list_entries_is_in_walk_order(seed):
    """m2k_list_entries (GetFilesByName / GetFilesByExt, utils.go:47-120) sorts a subset of entries
    into Walk order from the directory numbering alone; checked against the plain recursive walk."""
    import random
    rng = random.Random(seed)
    tp = lib.synth_tree_host(small_synth(seed=seed, n_files=3000, n_dirs=400, total=1 << 20), host_bytes=False)
    try:
        a = _tree_arrays(tp)
        ne = a["n"][1]
        order = _walk_order(a)
        assert sorted(order) == list(range(ne))
        for frac in (0.002, 0.05, 0.5, 1.0):
            feat = (C.c_uint32 * ne)()
            chosen = set()
            for e in range(ne):
                if rng.random() < frac:
                    feat[e] = 1
                    chosen.add(e)
            got = lib.list_entries(tp, 0, feat)
            assert got == [e for e in order if e in chosen]
    finally:
        lib.tree_free_host(tp)


def test_host_prepare_probe_runs_without_a_device():
    tp = lib.synth_tree_host(small_synth(seed=5, n_files=2000, n_dirs=300, total=1 << 20), host_bytes=False)
    try:
        ms = (C.c_double * 6)()
        nf, nr = C.c_uint32(), C.c_uint32()
        assert lib.load().m2k_host_prepare_probe(tp, None, 2, ms, C.byref(nf), C.byref(nr)) == 0
        assert nf.value >= 1 and nr.value >= 1 and all(0 <= v < 1e4 for v in ms)
    finally:
        lib.tree_free_host(tp)


@pytest.mark.parametrize("name", ["nasty", "samples", "multiignores", "random007"])
def test_list_entries_walk_order_on_packed_trees(tmp_path, name):
    """Same check on trees packed from the filesystem (symlinks, dot names, directories that sort
    between files)."""
    import random
    spec = next(g for g in GOLDENS if g["name"] == name)
    root = os.fsencode(treespec.materialize(spec, str(tmp_path / name)))
    tp = lib.pack_tree_host(root)
    try:
        a = _tree_arrays(tp)
        ne = a["n"][1]
        order = _walk_order(a)
        rng = random.Random(11)
        for frac in (0.3, 1.0):
            feat = (C.c_uint32 * max(ne, 1))()
            chosen = {e for e in range(ne) if rng.random() < frac}
            for e in chosen:
                feat[e] = 1
            assert lib.list_entries(tp, 0, feat) == [e for e in order if e in chosen]
    finally:
        lib.tree_free_host(tp)


def test_plan_shards_cuts_only_large_files():
    """m2k_plan_shards / m2k_chunk_locate: near-equal byte ranges whose cut points
    fall inside a file only when the file is large enough to be worth spreading over GPUs."""
    p = small_synth(seed=9, n_files=300, n_dirs=30, total=24 << 20, n_huge=2, huge_bytes=6 << 20)
    tp = lib.synth_tree_host(p, host_bytes=False)
    try:
        t = tp.contents
        lens = t.file_len[0:t.n_files]
        chunks = [(n + 65535) // 65536 for n in lens]
        total = sum(chunks)
        assert lib.chunk_count(tp) == total
        starts = [0]
        for c in chunks:
            starts.append(starts[-1] + c)
        for world in (1, 2, 3, 5, 8):
            for split_min in (1 << 20, 1 << 40):
                cuts = lib.plan_shards(tp, world, split_min)
                assert cuts[0] == 0 and cuts[-1] == total and cuts == sorted(cuts) and len(cuts) == world + 1
                for c in cuts[1:-1]:
                    if c in (0, total):
                        continue
                    f, k = lib.chunk_locate(tp, c)
                    assert starts[f] + k == c and 0 <= k < chunks[f]
                    if k:
                        assert lens[f] >= split_min  # a cut inside a file: only a large one
                if split_min == 1 << 20 and world > 1:
                    # balanced by cost (bytes + a fixed 4 KiB per chunk) to within a small file
                    cost = [0]
                    for n, c in zip(lens, chunks):
                        cost.append(cost[-1] + n + 4096 * c)
                    def cost_at(chunk):
                        if chunk == total:
                            return cost[-1]
                        f, k = lib.chunk_locate(tp, chunk)
                        return cost[f] + k * (65536 + 4096)
                    sizes = [cost_at(cuts[i + 1]) - cost_at(cuts[i]) for i in range(world)]
                    assert max(sizes) - min(sizes) <= 2 * ((1 << 20) + 4096 * 17)
    finally:
        lib.tree_free_host(tp)


def test_packer_reproduces_a_materialised_generator_tree(tmp_path):
    """A generator tree written to disk and packed again (threaded walk, per-directory reads, one file
    large enough to be read in pread segments) is the generator's tree, array for array and byte for byte."""
    from move2kube_b200 import synthfs
    p = small_synth(seed=17, n_files=4000, n_dirs=300, total=70 << 20, n_huge=1, huge_bytes=40 << 20)
    src = lib.synth_tree_host(p, host_bytes=True)
    root = os.fsencode(str(tmp_path / "tree"))
    try:
        n = synthfs.materialize(src, root)
        assert n == src.contents.content_bytes
        got = lib.pack_tree_host(root)
        try:
            a, b = _tree_arrays(src), _tree_arrays(got)
            assert max(src.contents.file_len[0:src.contents.n_files]) > 32 << 20
            for k in a:
                assert a[k] == b[k], k
        finally:
            lib.tree_free_host(got)
    finally:
        lib.tree_free_host(src)


def test_host_thread_team_is_shared_safely_between_caller_threads():
    """Contexts driven from different host threads share one team of worker threads: two callers at
    once (ctypes drops the GIL during the call) must both get complete, identical answers."""
    import threading
    p = _abi.SynthParams(seed=23, n_files=30000, n_dirs=3000, max_depth=6, n_ignore_files=800, total_bytes=1 << 20)
    tp = lib.synth_tree_host(p, host_bytes=False)
    try:
        L = lib.load()
        out = {}

        def work(k):
            ms = (C.c_double * 6)()
            nf, nr = C.c_uint32(), C.c_uint32()
            rc = L.m2k_host_prepare_probe(tp, None, 20, ms, C.byref(nf), C.byref(nr))
            out[k] = (rc, nf.value, nr.value)

        th = [threading.Thread(target=work, args=(k,)) for k in range(4)]
        for x in th:
            x.start()
        for x in th:
            x.join()
        assert len(out) == 4 and len(set(out.values())) == 1 and out[0][0] == 0 and out[0][2] > 4000
    finally:
        lib.tree_free_host(tp)


def _build_c_example(out_dir):
    """examples/detect.c against include/m2kdetect.h and libm2kdetect.so: the C ABI used from C, no binding layer"""
    exe = os.path.join(str(out_dir), "detect")
    libdir = os.path.join(REPO, "move2kube_b200")
    subprocess.check_call(["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(REPO, "include"), os.path.join(REPO, "examples", "detect.c"),
                           "-L", libdir, "-lm2kdetect", "-Wl,-rpath," + libdir, "-o", exe])
    return exe


def test_c_example_builds_links_and_fails_loudly_without_a_device(tmp_path):
    import torch
    exe = _build_c_example(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a device is present: tests/test_gpu_mirror.py runs the example")
    p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
    assert p.returncode == 1 and p.stdout == "" and "no CPU path" in p.stderr
