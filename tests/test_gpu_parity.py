"""GPU parity tests (pytest -m gpu): the CUDA path, called through the C ABI, against the golden
fixtures and the CPU oracle on the same inputs.  Bit-exact: everything here is integer/byte work."""
import ctypes as C
import os

import pytest

import golden_check as gc
import treespec
from dockerfile_kats import KATS, reference_corpus
from move2kube_b200 import _abi, lib

pytestmark = pytest.mark.gpu

GOLDENS = gc.load_goldens()


@pytest.fixture(scope="module")
def ctx():
    c = lib.Context(0)
    yield c
    c.close()


def compare_with_oracle(oracle, tree, res, ores, content=True):
    """Every array the C ABI returns against the oracle's, for every directory and file."""
    t, r, o = tree.contents, res.contents, ores.contents
    nd, nf = t.n_dirs, t.n_files
    assert (r.n_dirs, r.n_files) == (nd, nf)
    assert r.dir_opts[0:nd] == o.dir_opts[0:nd]
    assert r.dir_state[0:nd] == o.dir_state[0:nd]
    assert r.service_dirs[0:r.n_services] == o.service_dirs[0:o.n_services]
    # the oracle fills the ignore flags of visited directories only
    for d in range(nd):
        if o.dir_state[d] not in (_abi.DIR_NOT_VISITED, _abi.DIR_WALK_ERROR):
            assert r.dir_ignore[d] == o.dir_ignore[d], d
    assert [r.ignore_dirs[i] for i in range(r.n_ignore_dirs)] == [o.ignore_dirs[i] for i in range(o.n_ignore_dirs)]
    assert [r.ignore_contents[i] for i in range(r.n_ignore_contents)] == \
        [o.ignore_contents[i] for i in range(o.n_ignore_contents)]
    if content:
        assert r.file_flags[0:nf] == o.file_flags[0:nf]
        assert r.dir_main_file[0:nd] == o.dir_main_file[0:nd]


@pytest.mark.parametrize("spec", [g for g in GOLDENS if not g.get("nonexistent")], ids=lambda g: g["name"])
@pytest.mark.parametrize("flavor", ["bash", "dash"])
def test_golden_fixture(ctx, oracle, tmp_path, spec, flavor):
    root = os.fsencode(treespec.materialize(spec, str(tmp_path / spec["name"])))
    fl = _abi.SH_BASH if flavor == "bash" else _abi.SH_DASH
    pre = [oracle.join(root, os.fsencode(p)) if p != "." else root for p in spec.get("pre", [])]
    tree = ctx.pack_tree(root)
    try:
        res = ctx.detect(tree, fl, pre)
        try:
            assert res.contents.n_kernel_launches >= 3
            if not gc.has_user_scripts(spec):
                gc.check(spec, flavor, gc.TreeView(tree), res, root)
                n_out = gc.check_outputs(spec, flavor, gc.TreeView(tree), tree, res)
                assert n_out >= sum(len(v) for v in spec["expect"][flavor]["outputs"].values())
            ores = oracle.detect(tree, fl, pre)
            compare_with_oracle(oracle, tree, res, ores)
            oracle.lib().m2ko_result_free(ores)
        finally:
            ctx.result_free(res)
    finally:
        ctx.tree_free(tree)


def test_nonexistent_root_is_an_io_error(ctx, tmp_path):
    with pytest.raises(lib.M2KError) as ei:
        ctx.pack_tree(os.fsencode(str(tmp_path / "missing")))
    assert ei.value.code == _abi.ERR_IO


def test_dockerfiles_that_ship_in_the_reference_tree(ctx, oracle, tmp_path):
    """The reference tree's own 16 Dockerfiles and 31 of its other files, laid out as in the reference, through
    the GPU path: flags equal the oracle's and the expectation made by inspection (make_dockerfile_corpus.py)."""
    root = tmp_path / "ref"
    corpus = reference_corpus()
    for path, data, want in corpus:
        (root / path).parent.mkdir(parents=True, exist_ok=True)
        (root / path).write_bytes(data)
    tree = ctx.pack_tree(os.fsencode(str(root)))
    res = ctx.detect(tree)
    view = gc.TreeView(tree)
    got = {os.fsdecode(view.file_rel[f]): res.contents.file_flags[f] for f in range(view.n_files)}
    for path, data, want in corpus:
        assert got[path] == oracle.file_flags(data), path
        assert bool(got[path] & _abi.FILE_DOCKERFILE) == want, path
    ctx.result_free(res)
    ctx.tree_free(tree)


def test_dockerfile_known_answers(ctx, oracle, tmp_path):
    root = tmp_path / "kats"
    root.mkdir()
    for i, (name, data, want) in enumerate(KATS):
        (root / ("%03d_%s" % (i, name))).write_bytes(data)
    broot = os.fsencode(str(root))
    tree = ctx.pack_tree(broot)
    res = ctx.detect(tree)
    view = gc.TreeView(tree)
    got = {}
    for f in range(view.n_files):
        got[os.fsdecode(view.file_rel[f])] = res.contents.file_flags[f]
    for i, (name, data, want) in enumerate(KATS):
        fl = got["%03d_%s" % (i, name)]
        assert fl == oracle.file_flags(data), name
        assert bool(fl & _abi.FILE_DOCKERFILE) == want, name
    ctx.result_free(res)
    ctx.tree_free(tree)


def test_dockerfile_known_answers_through_the_striped_path(ctx, oracle, tmp_path):
    """The known-answer vectors followed by 12 KiB of further lines (after a blank-line separator, and
    glued on directly so that trailing continuations run into them): bodies of a dozen stripes, with
    the vector's own lines in the first one.  GPU against the oracle on identical bytes."""
    root = tmp_path / "kats_long"
    root.mkdir()
    pad = b"RUN echo a line of padding that is valid in any Dockerfile body\n" * 200
    bad = b"RUN true\n" * 1400 + b"ENV missingvalue\n"
    datas = {}
    for i, (name, data, want) in enumerate(KATS):
        datas["%03d_a_%s" % (i, name)] = data + pad
        datas["%03d_b_%s" % (i, name)] = data + b"\n" + pad
        datas["%03d_c_%s" % (i, name)] = data + b"\n" + bad
    for k, v in datas.items():
        (root / k).write_bytes(v)
    tree = ctx.pack_tree(os.fsencode(str(root)))
    res = ctx.detect(tree)
    view = gc.TreeView(tree)
    n_df = 0
    for f in range(view.n_files):
        name = os.fsdecode(view.file_rel[f])
        fl = res.contents.file_flags[f]
        assert fl == oracle.file_flags(datas[name]), name
        n_df += bool(fl & _abi.FILE_DOCKERFILE)
    assert n_df >= 20
    ctx.result_free(res)
    ctx.tree_free(tree)


def _adversarial_files():
    """Content cases around the kernel's tiling: 16-byte units, 512-byte rows, 64 KiB chunks."""
    out = {}
    for pos in [0, 1, 2, 3, 5, 8, 9, 13, 15, 16, 505, 508, 511, 512, 4089, 65528, 65529, 65531, 65535, 65536, 131064, 131071]:
        body = bytearray(b"x" * (pos + 8 + 40))
        for k in range(60, len(body), 61):
            body[k] = 0x0A
        body[pos:pos + 8] = b"__main__"
        out["main_at_%d.py" % pos] = bytes(body)
        out["nomain_at_%d.py" % pos] = bytes(body[:pos + 7])  # cut one byte short
    for pat in [b"__main_", b"_main__", b"__ma", b"main", b"ain_", b"_mai", b"__mainx__main_", b"__mai\nn__"]:
        out["near_%d" % len(out)] = b"abc " * 7 + pat + b" tail\n"
    out["exact8"] = b"__main__"
    out["seven"] = b"__main_"
    out["empty"] = b""
    # line lengths around bufio.MaxScanTokenSize, at every alignment that matters
    for lead in [0, 1, 15, 16, 17, 511, 65535, 65536, 65537]:
        for ln in [65534, 65535, 65536, 65537]:
            for tail in [b"", b"\n", b"\r\n", b"\nshort\n"]:
                key = "line_%d_%d_%d" % (lead, ln, len(tail))
                out[key] = (b"a" * lead + b"\n" if lead else b"") + b"b" * ln + tail
    out["long_first_of_three_chunks"] = b"q" * 200000
    out["long_between"] = b"short\n" * 100 + b"L" * 70000 + b"\n" + b"short\n" * 100
    out["nl_every_unit"] = (b"123456789012345\n") * 9000
    out["nl_sparse_ok"] = (b"z" * 65535 + b"\n") * 4
    out["nl_sparse_bad"] = (b"z" * 65535 + b"\n") * 2 + b"z" * 65536 + b"\n" + b"z" * 100
    return out


def test_adversarial_content(ctx, oracle, tmp_path):
    root = tmp_path / "adv"
    root.mkdir()
    files = _adversarial_files()
    for name, data in files.items():
        (root / name).write_bytes(data)
    tree = ctx.pack_tree(os.fsencode(str(root)))
    res = ctx.detect(tree)
    view = gc.TreeView(tree)
    bad = []
    for f in range(view.n_files):
        name = os.fsdecode(view.file_rel[f])
        want = oracle.file_flags(files[name])
        if res.contents.file_flags[f] != want:
            bad.append((name, res.contents.file_flags[f], want))
    assert not bad, bad[:10]
    ctx.result_free(res)
    ctx.tree_free(tree)


def synth_params(seed, n_files, n_dirs, total, depth=6, n_ignore=6, **kw):
    return _abi.SynthParams(seed=seed, n_files=n_files, n_dirs=n_dirs, max_depth=depth, n_ignore_files=n_ignore,
                            total_bytes=total, n_huge=kw.get("n_huge", 0), huge_bytes=kw.get("huge_bytes", 0),
                            tiny_min=kw.get("tiny_min", 0), tiny_max=kw.get("tiny_max", 0))


@pytest.mark.parametrize("seed,n_files,n_dirs,total", [(7, 4000, 200, 16 << 20), (3, 1500, 120, 96 << 20)])
def test_synthetic_tree_vs_oracle(ctx, oracle, seed, n_files, n_dirs, total):
    tree = ctx.synth_tree(synth_params(seed, n_files, n_dirs, total), host_bytes=True)
    try:
        res = ctx.detect(tree)
        ores = oracle.detect(tree, n_threads=8)
        compare_with_oracle(oracle, tree, res, ores)
        r = res.contents
        assert r.n_services > 0
        flags = r.file_flags[0:r.n_files]
        assert any(x & _abi.FILE_DOCKERFILE for x in flags) and any(x & _abi.FILE_HAS_MAIN for x in flags)
        oracle.lib().m2ko_result_free(ores)
        ctx.result_free(res)
    finally:
        ctx.tree_free(tree)


def test_device_fill_equals_host_fill_and_staged_equals_whole(ctx, oracle):
    """m2k_synth_fill_device writes the bytes m2k_synth_fill_host writes, and upload/scan/finalize
    returns what m2k_detect returns."""
    p = synth_params(11, 3000, 150, 48 << 20)
    host_tree = ctx.synth_tree(p, host_bytes=True)
    dev_tree = ctx.synth_tree(p, host_bytes=False)
    try:
        whole = ctx.detect(host_tree)
        dt = ctx.upload(dev_tree)
        ctx.synth_fill_device(dt, 11)
        ctx.scan(dt)
        assert ctx.scan_kernel_ms() > 0
        staged = ctx.finalize(dt)
        a, b = whole.contents, staged.contents
        nd, nf = a.n_dirs, a.n_files
        for fld in ("dir_opts", "dir_subtree", "dir_main_file", "dir_state", "dir_ignore"):
            assert getattr(a, fld)[0:nd] == getattr(b, fld)[0:nd], fld
        assert a.file_flags[0:nf] == b.file_flags[0:nf]
        assert a.service_dirs[0:a.n_services] == b.service_dirs[0:b.n_services]
        ctx.devtree_free(dt)
        ctx.result_free(staged)
        ctx.result_free(whole)
    finally:
        ctx.tree_free(host_tree)
        ctx.tree_free(dev_tree)


def test_sharded_scan_merges_to_the_whole(ctx, oracle):
    """Two shards of the files scanned separately (as two ranks would), partials merged with the
    gather kernel, finalised: same answer as one pass.  Emulates the N>1 path on one GPU."""
    import torch
    p = synth_params(5, 2500, 140, 40 << 20)
    tree = ctx.synth_tree(p, host_bytes=True)
    try:
        whole = ctx.detect(tree)
        nf, nd = tree.contents.n_files, tree.contents.n_dirs
        mid = nf // 2
        parts = []
        for a, b in ((0, mid), (mid, nf)):
            dt = ctx.upload(tree, a, b)
            ctx.scan(dt)
            parts.append(dt)
        allc = torch.empty((2, nd), dtype=torch.int32, device="cuda")
        allm = torch.empty((2, nd), dtype=torch.int32, device="cuda")
        for i, dt in enumerate(parts):
            ctx.partials(dt, allc[i].data_ptr(), allm[i].data_ptr())
        torch.cuda.synchronize()
        results = []
        for i, dt in enumerate(parts):
            ctx.merge_partials(dt, allc.data_ptr(), allm.data_ptr(), 2)
            results.append(ctx.finalize(dt))
        w = whole.contents
        for i, (a, b) in enumerate(((0, mid), (mid, nf))):
            r = results[i].contents
            assert r.dir_state[0:nd] == w.dir_state[0:nd]
            assert r.dir_opts[0:nd] == w.dir_opts[0:nd]
            assert r.dir_main_file[0:nd] == w.dir_main_file[0:nd]
            assert r.dir_subtree[0:nd] == w.dir_subtree[0:nd]
            assert r.file_flags[a:b] == w.file_flags[a:b]
        for i, dt in enumerate(parts):
            ctx.result_free(results[i])
            ctx.devtree_free(dt)
        ctx.result_free(whole)
    finally:
        ctx.tree_free(tree)


def test_names_only_mode(ctx, oracle, tmp_path):
    spec = next(g for g in GOLDENS if g["name"] == "nasty")
    root = os.fsencode(treespec.materialize(spec, str(tmp_path / "nasty")))
    tree = ctx.pack_tree(root, read_content=False)
    res = ctx.detect(tree, skip_content=True)
    gc_spec = dict(spec)
    gc_spec["expect"] = {"bash": dict(spec["expect"]["bash"], main_script={})}
    gc.check(gc_spec, "bash", gc.TreeView(tree), res, root)
    assert not any(res.contents.file_flags[0:res.contents.n_files])
    ctx.result_free(res)
    ctx.tree_free(tree)


def test_very_deep_tree(ctx, oracle, tmp_path):
    """Paths deeper than the walk kernel's in-register ancestor chain (64) take its slow branch."""
    root = tmp_path / "deep"
    p = root
    rel = []
    for i in range(90):
        name = "d%d" % (i % 7)
        p = p / name
        rel.append(name)
    p.mkdir(parents=True)
    (root.joinpath(*rel[:83]) / "package.json").write_text("{}")
    (root / ".m2kignore").write_text(".\n" + "/".join(rel[:70]) + "/\n" + "/".join(rel[:75]).upper() + "/\n"
                                     + "/".join(rel[:86]) + "/*\n")
    broot = os.fsencode(str(root))
    tree = ctx.pack_tree(broot)
    res = ctx.detect(tree)
    ores = oracle.detect(tree)
    compare_with_oracle(oracle, tree, res, ores)
    r = res.contents
    states = r.dir_state[0:r.n_dirs]
    assert states.count(_abi.DIR_IGNORED_DESCEND) == 3 and r.n_services == 1 and _abi.DIR_NOT_VISITED in states
    oracle.lib().m2ko_result_free(ores)
    ctx.result_free(res)
    ctx.tree_free(tree)


# ---- shards cut through files -------------------------------------------
def run_chunk_sharded(ctx, tree, cuts, fill_seed=None):
    """Emulates an N-rank pass on one GPU: every shard [cuts[r], cuts[r+1]) is uploaded and scanned on its
    own, partials / file flags / pieces are exchanged the way the ranks exchange them (rank order),
    and shard 0 merges, settles and finalises.  Returns (result, devtrees, settled files)."""
    import torch
    from move2kube_b200 import multi
    n = len(cuts) - 1
    nd, nf = tree.contents.n_dirs, tree.contents.n_files
    dts = []
    for r in range(n):
        dt = ctx.upload_chunks(tree, cuts[r], cuts[r + 1])
        if fill_seed is not None:
            ctx.synth_fill_device(dt, fill_seed)
        ctx.scan(dt)
        dts.append(dt)
    allc = torch.empty((n, nd), dtype=torch.int32, device="cuda")
    allm = torch.empty((n, nd), dtype=torch.int32, device="cuda")
    pieces = torch.empty((n, 64), dtype=torch.uint8, device="cuda")
    flags = torch.zeros((nf,), dtype=torch.uint8, device="cuda")
    one = torch.empty((nf,), dtype=torch.uint8, device="cuda")
    for r, dt in enumerate(dts):
        ctx.partials(dt, allc[r].data_ptr(), allm[r].data_ptr())
        ctx.pieces(dt, pieces[r].data_ptr())
        ctx.file_flags_get(dt, one.data_ptr())
        torch.cuda.synchronize()
        flags += one
    torch.cuda.synchronize()
    ctx.merge_partials(dts[0], allc.data_ptr(), allm.data_ptr(), n)
    ctx.file_flags_set(dts[0], flags.data_ptr())
    ctx.merge_pieces(dts[0], pieces.data_ptr(), n)
    settled = multi.settle_undecided(ctx, tree, dts[0], multi.split_files(tree, cuts), fill_seed)
    return ctx.finalize(dts[0]), dts, settled


def run_chunk_sharded_peer(ctx, tree, cuts, fill_seed=None):
    """The same N-rank pass with the exchange over peer memory (m2k_exchange_*): every rank has its own
    context and stream (all on this one GPU: the root's arrays are then simply local memory, the
    protocol is the one real peers run), ranks 1.. reduce straight into rank 0's arrays from inside
    their per-file kernels, rank 0 collects, settles and finalises."""
    from move2kube_b200 import multi
    n = len(cuts) - 1
    ctxs = [ctx] + [lib.Context(0) for _ in range(n - 1)]
    dts = []
    for r in range(n):
        dt = ctxs[r].upload_chunks(tree, cuts[r], cuts[r + 1])
        if fill_seed is not None:
            ctxs[r].synth_fill_device(dt, fill_seed)
        dts.append(dt)
    ctx.exchange_open(dts[0], n)
    for r in range(1, n):
        ctxs[r].exchange_attach_local(dts[r], r, ctx, dts[0])
    res = None
    for _ in range(2):  # twice: the second pass reuses the mailbox and every buffer of the first
        if res is not None:
            ctx.result_free(res)
        for r in reversed(range(n)):  # the root last: the others must wait for its "arrays cleared" flag
            ctxs[r].scan(dts[r])
        ctx.exchange_collect(dts[0])
        for r in range(1, n):
            ctxs[r].exchange_wait(dts[r])
        settled = multi.settle_undecided(ctx, tree, dts[0], multi.split_files(tree, cuts), fill_seed)
        res = ctx.finalize(dts[0])
    for r in range(1, n):
        ctxs[r].devtree_free(dts[r])
        ctxs[r].close()
    return res, dts[:1], settled


def sharded_runner(kind):
    return run_chunk_sharded if kind == "gather" else run_chunk_sharded_peer


def assert_same_result(tree, got, want):
    t, g, w = tree.contents, got.contents, want.contents
    nd, nf = t.n_dirs, t.n_files
    assert g.file_flags[0:nf] == w.file_flags[0:nf]
    for name in ("dir_state", "dir_opts", "dir_main_file", "dir_subtree", "dir_direct", "dir_ignore"):
        assert getattr(g, name)[0:nd] == getattr(w, name)[0:nd], name
    assert g.service_dirs[0:g.n_services] == w.service_dirs[0:w.n_services]


@pytest.mark.parametrize("exchange", ["gather", "peer"])
@pytest.mark.parametrize("world", [2, 3, 5])
def test_chunk_shards_of_a_synthetic_tree_merge_to_the_whole(ctx, oracle, world, exchange):
    """Multi-MiB files spread over several shards, bytes generated on the device shard by shard."""
    p = _abi.SynthParams(seed=21, n_files=1500, n_dirs=90, max_depth=5, n_ignore_files=3, total_bytes=60 << 20,
                         n_huge=3, huge_bytes=9 << 20, n_large=2, large_bytes=3 << 20)
    tree = ctx.synth_tree(p, host_bytes=True)
    try:
        whole = ctx.detect(tree)
        ores = oracle.detect(tree)
        compare_with_oracle(oracle, tree, whole, ores)
        oracle.lib().m2ko_result_free(ores)
        from move2kube_b200 import multi
        t = tree.contents
        chunks = [(n + 65535) // 65536 for n in t.file_len[0:t.n_files]]
        starts = [0]
        for c in chunks:
            starts.append(starts[-1] + c)
        total = starts[-1]
        big = [f for f in range(t.n_files) if t.file_len[f] >= 1 << 20]
        assert len(big) >= 5
        plans = [lib.plan_shards(tree, world, 1 << 20)]
        if world == 2:   # one cut in the middle of each large file
            plans += [[0, starts[f] + chunks[f] // 2, total] for f in big]
        if world == 3:   # both cuts inside the same large file: the middle shard is one piece of it
            plans += [[0, starts[f] + chunks[f] // 3, starts[f] + 2 * chunks[f] // 3, total] for f in big[:2]]
        n_split = 0
        for cuts in plans:
            n_split += len(multi.split_files(tree, cuts))
            for seed in (None, p.seed):  # bytes uploaded from the host / generated in HBM
                res, dts, _ = sharded_runner(exchange)(ctx, tree, cuts, fill_seed=seed)
                assert_same_result(tree, res, whole)
                ctx.result_free(res)
                for dt in dts:
                    ctx.devtree_free(dt)
        assert world == 5 or n_split
        ctx.result_free(whole)
    finally:
        ctx.tree_free(tree)


def _big_file_tree(root):
    """Files of a few 64 KiB chunks whose interesting bytes sit exactly on chunk boundaries."""
    CH = 65536
    os.makedirs(root / "app")
    line = b"x = 1  # padding line of text\n"
    def text(n):
        return (line * (n // len(line) + 1))[:n - 1] + b"\n"
    # "__main__" straddling the boundary between chunks 1 and 2 at every offset
    for k in range(1, 8):
        body = bytearray(text(5 * CH - 100))
        body[2 * CH - k:2 * CH - k + 8] = b"__main__"
        (root / "app" / ("m%d.py" % k)).write_bytes(bytes(body))
    (root / "app" / "requirements.txt").write_bytes(b"flask\n")
    # a line of exactly 65536 / 65535 bytes straddling a boundary
    for name, n in (("long.txt", 65536), ("short.txt", 65535)):
        (root / name).write_bytes(text(CH + 1000) + b"y" * n + b"\n" + text(2 * CH))
    # one very long line covering several whole chunks
    (root / "huge_line.bin").write_bytes(text(CH // 2) + b"z" * (3 * CH) + b"\n" + text(CH))
    # no newline at all
    (root / "nonl.bin").write_bytes(b"q" * (2 * CH + 17))
    # a large valid Dockerfile, a large invalid one (error deep in the body), one whose head is longer than a chunk
    run = b"RUN echo hello && \\\n    echo world\n"
    (root / "Dockerfile.big").write_bytes(b"FROM alpine:3 AS base\n" + run * (3 * CH // len(run)))
    (root / "Dockerfile.bad").write_bytes(b"FROM alpine:3\n" + run * (2 * CH // len(run)) + b"ENV onlykey\n" + run * 100)
    (root / "Dockerfile.comments").write_bytes(b"# a comment line that is not a directive\n" * (CH // 20) + b"ARG V=1\nFROM alpine:$V\n" + run * 50)
    (root / "notes.big").write_bytes(b"hello world, this is not a Dockerfile\n" * (4 * CH // 38))
    (root / "blob.big").write_bytes(bytes([0x89, 0x50, 0x4E, 0x47, 0x0D, 0x0A]) + bytes((i * 7 + 3) % 251 for i in range(3 * CH)))
    (root / "empty").write_bytes(b"")
    return root


@pytest.mark.parametrize("exchange", ["gather", "peer"])
def test_every_cut_through_files_with_boundary_cases(ctx, oracle, tmp_path, exchange):
    """Two shards with the cut at every chunk of the tree, and a few three-way cuts: "__main__", long
    lines and Dockerfile heads that straddle the cut come out as in one pass."""
    root = os.fsencode(str(_big_file_tree(tmp_path / "big")))
    tree = ctx.pack_tree(root)
    try:
        whole = ctx.detect(tree)
        ores = oracle.detect(tree)
        compare_with_oracle(oracle, tree, whole, ores)
        oracle.lib().m2ko_result_free(ores)
        w = whole.contents
        view = gc.TreeView(tree)
        names = {os.path.basename(view.file_rel[f]): w.file_flags[f] for f in range(tree.contents.n_files)}
        assert names[b"Dockerfile.big"] & _abi.FILE_DOCKERFILE and names[b"Dockerfile.comments"] & _abi.FILE_DOCKERFILE
        assert not names[b"Dockerfile.bad"] & _abi.FILE_DOCKERFILE and names[b"Dockerfile.bad"] & _abi.FILE_FROM_FIRST
        assert names[b"long.txt"] & _abi.FILE_LONG_LINE and not names[b"short.txt"] & _abi.FILE_LONG_LINE
        assert all(names[b"m%d.py" % k] & _abi.FILE_HAS_MAIN for k in range(1, 8))
        total = lib.chunk_count(tree)
        plans = [[0, c, total] for c in range(1, total)] + [[0, total // 3, 2 * total // 3, total], [0, 1, total - 1, total],
                                                            [0, total // 2, total // 2, total]]
        settled_any = False
        for cuts in plans:
            res, dts, settled = sharded_runner(exchange)(ctx, tree, cuts)
            settled_any = settled_any or bool(settled)
            assert_same_result(tree, res, whole)
            ctx.result_free(res)
            for dt in dts:
                ctx.devtree_free(dt)
        assert settled_any  # the large Dockerfiles had to be settled by a pass of their own
        ctx.result_free(whole)
    finally:
        ctx.tree_free(tree)


@pytest.mark.parametrize("n", [2, 3])
def test_detect_multi_from_one_process(ctx, oracle, n):
    """m2k_detect_multi: one host thread per context, shards reduced into the first context's arrays over
    peer memory.  With fewer GPUs than contexts the extra contexts share device 0 (same protocol, local
    memory); with enough GPUs every context gets its own device.  Files of 70 MiB (nearly all of the bytes) are cut between shards."""
    import torch
    n_gpus = torch.cuda.device_count()
    ctxs = [ctx] + [lib.Context(r if r < n_gpus else 0) for r in range(1, n)]
    p = _abi.SynthParams(seed=44, n_files=1200, n_dirs=80, max_depth=5, n_ignore_files=3, total_bytes=8 << 20,
                         n_huge=4, huge_bytes=70 << 20, n_large=2, large_bytes=3 << 20)
    tree = ctx.synth_tree(p, host_bytes=True)
    try:
        from move2kube_b200 import multi
        cuts = lib.plan_shards(tree, n, 64 << 20, None)
        assert multi.split_files(tree, cuts), "the test tree should have a file cut between shards"
        whole = ctx.detect(tree)
        ores = oracle.detect(tree, n_threads=4)
        compare_with_oracle(oracle, tree, whole, ores)
        for _ in range(2):
            res = lib.detect_multi(ctxs, tree)
            assert_same_result(tree, res, whole)
            compare_with_oracle(oracle, tree, res, ores)
            ctx.result_free(res)
        oracle.lib().m2ko_result_free(ores)
        ctx.result_free(whole)
    finally:
        ctx.tree_free(tree)
        for c in ctxs[1:]:
            c.close()


def test_pageable_tree_goes_through_the_staging_blocks(ctx, oracle):
    """A tree whose bytes sit in ordinary (pageable) host memory is streamed to the device through the
    context's two 32 MiB pinned blocks; 90 MB makes the blocks wrap around.  Same result as the oracle."""
    p = synth_params(13, 2500, 150, 90 << 20)
    tree = lib.synth_tree_host(p, host_bytes=True)  # no context: plain malloc
    try:
        for _ in range(2):
            res = ctx.detect(tree)
            ores = oracle.detect(tree)
            compare_with_oracle(oracle, tree, res, ores)
            oracle.lib().m2ko_result_free(ores)
            ctx.result_free(res)
    finally:
        lib.tree_free_host(tree)


@pytest.mark.parametrize("name", ["samples", "nasty", "multiignores", "ignorecontents", "emptydir", "random007", "random019"])
def test_pack_detect_streams_the_same_answer(ctx, tmp_path, name):
    """m2k_pack_detect (bytes streamed from the files through the pinned blocks, never gathered on the
    host) against m2k_pack_tree + m2k_detect on the same directory: every array."""
    spec = next(g for g in GOLDENS if g["name"] == name)
    root = os.fsencode(treespec.materialize(spec, str(tmp_path / name)))
    tree = ctx.pack_tree(root)
    want = ctx.detect(tree)
    for skip in (False, True):
        if skip:
            ctx.result_free(want)
            want = ctx.detect(tree, skip_content=True)
        t2, got = ctx.pack_detect(root, skip_content=skip)
        assert not t2.contents.bytes
        # (names only: the streamed tree then carries no sizes except for the rule files)
        assert _meta(t2)[:7] == _meta(tree)[:7] and (skip or _meta(t2) == _meta(tree))
        assert_same_result(tree, got, want)
        g, w = got.contents, want.contents
        assert [g.ignore_dirs[i] for i in range(g.n_ignore_dirs)] == [w.ignore_dirs[i] for i in range(w.n_ignore_dirs)]
        assert [g.ignore_contents[i] for i in range(g.n_ignore_contents)] == [w.ignore_contents[i] for i in range(w.n_ignore_contents)]
        ctx.result_free(got)
        ctx.tree_free(t2)
    ctx.result_free(want)
    ctx.tree_free(tree)


def _meta(tp):
    t = tp.contents
    return (t.n_dirs, t.n_entries, t.n_files, t.dir_parent[0:t.n_dirs], t.entry_dir[0:t.n_entries], t.entry_type[0:t.n_entries],
            bytes(bytearray(t.names[0:t.n_name_bytes])), t.file_len[0:t.n_files] if t.n_files else [])


def test_pack_detect_on_a_large_materialised_tree(ctx, oracle, tmp_path):
    """Enough bytes for the two pinned blocks to wrap around several times (150 MB), one file larger
    than a block: the streamed pass equals the oracle's answer on the generator's tree."""
    from move2kube_b200 import synthfs
    p = synth_params(29, 3000, 200, 150 << 20, n_huge=1, huge_bytes=70 << 20)
    src = lib.synth_tree_host(p, host_bytes=True)
    root = os.fsencode(str(tmp_path / "tree"))
    try:
        synthfs.materialize(src, root)
        ores = oracle.detect(src)
        t2, got = ctx.pack_detect(root)
        o, g = ores.contents, got.contents
        nd, nf = src.contents.n_dirs, src.contents.n_files
        assert g.file_flags[0:nf] == o.file_flags[0:nf]
        assert g.dir_opts[0:nd] == o.dir_opts[0:nd] and g.dir_state[0:nd] == o.dir_state[0:nd]
        assert g.dir_main_file[0:nd] == o.dir_main_file[0:nd]
        oracle.lib().m2ko_result_free(ores)
        ctx.result_free(got)
        ctx.tree_free(t2)
    finally:
        lib.tree_free_host(src)
