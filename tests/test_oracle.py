"""CPU tests: the C oracle against the golden fixtures (made by the reference's own detect scripts)."""
import os

import pytest

import golden_check as gc
import treespec
from move2kube_b200 import _abi

GOLDENS = gc.load_goldens()


@pytest.mark.parametrize("spec", GOLDENS, ids=[g["name"] for g in GOLDENS])
@pytest.mark.parametrize("flavor", ["bash", "dash"])
def test_oracle_matches_golden(oracle, tmp_path, spec, flavor):
    if gc.has_user_scripts(spec):
        pytest.skip("user-supplied detect scripts are exercised at the host-mirror level")
    root = str(tmp_path / spec["name"])
    if spec.get("nonexistent"):
        rc, tree = oracle.pack_fs(os.fsencode(root))
        assert rc == _abi.ERR_IO
        assert spec["expect"][flavor]["services"] == []
        return
    treespec.materialize(spec, root)
    broot = os.fsencode(root)
    rc, tree = oracle.pack_fs(broot)
    assert rc == 0
    try:
        pre = [oracle.join(broot, os.fsencode(p)) if p != "." else broot for p in spec.get("pre", [])]
        res = oracle.detect(tree, _abi.SH_BASH if flavor == "bash" else _abi.SH_DASH, pre=pre)
        try:
            gc.check(spec, flavor, gc.TreeView(tree), res, broot)
        finally:
            oracle.lib().m2ko_result_free(res)
    finally:
        oracle.lib().m2ko_tree_free(tree)


@pytest.mark.skipif(not os.path.isdir("/root/reference/samples"), reason="reference tree not on this machine")
def test_oracle_on_the_real_samples_tree(oracle):
    """BASELINE.json configs[0]: the reference's own samples/, with the real file contents."""
    spec = next(g for g in GOLDENS if g["name"] == "samples")
    broot = b"/root/reference/samples"
    rc, tree = oracle.pack_fs(broot)
    assert rc == 0
    res = oracle.detect(tree)
    view = gc.TreeView(tree)
    gc.check(spec, "bash", view, res, broot)
    # the three files isDockerFile should accept (by inspection; unpinned)
    got = sorted(os.fsdecode(view.file_rel[f]) for f in range(view.n_files)
                 if res.contents.file_flags[f] & _abi.FILE_DOCKERFILE)
    assert got == ["docker-compose/api/Dockerfile", "docker-compose/web/Dockerfile", "dockerfile/Dockerfile"]
    oracle.lib().m2ko_result_free(res)
    oracle.lib().m2ko_tree_free(tree)


ASSETS_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "m2kassets")


@pytest.mark.skipif(not os.path.isdir(ASSETS_DIR), reason="oracle/_ref/m2kassets (the reference's scripts) not unpacked")
@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_c_oracle_agrees_with_the_reference_scripts_on_generator_trees(oracle, tmp_path, seed):
    """The bench generator's trees, materialised on disk: the reference's own detect scripts (driven by
    the restated Go control flow) and the plain-C oracle must take the same decision for every
    directory -- this is what lets the C oracle stand in for the scripts at bench scale (config B is
    checked the same way by oracle/script_baseline.py --workload B)."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(ASSETS_DIR), ".."))
    import script_baseline as sb
    from move2kube_b200 import _abi, lib
    p = _abi.SynthParams(seed=seed, n_files=500, n_dirs=80, max_depth=6, n_ignore_files=3, total_bytes=3_000_000)
    tp = lib.synth_tree_host(p, host_bytes=True)
    root = os.fsencode(str(tmp_path / "tree"))
    sb.materialize(tp, root)
    lib.tree_free_host(tp)
    _, services, err, trace, ign_d, ign_c = sb.script_pass(root)
    assert err is None and len(trace) >= 5
    _, tree, res = sb.c_oracle_pass(root)
    bad, lists_ok = sb.compare(root, trace, services, ign_d, ign_c, tree, res)
    assert bad == [] and lists_ok


@pytest.mark.parametrize("spec", [g for g in gc.load_goldens() if not g.get("nonexistent") and not gc.has_user_scripts(g)], ids=lambda g: g["name"])
@pytest.mark.parametrize("flavor", ["bash", "dash"])
def test_option_json_matches_what_the_scripts_print(oracle, tmp_path, spec, flavor):
    """f3 without a GPU: m2k_option_json (host code of the product library) over the ORACLE's per-directory arrays
    must reproduce the stdout the reference's scripts printed for every accepted (directory, option) of every
    golden tree.  The -m gpu tier runs the same check over the device's arrays (test_gpu_parity.test_golden_fixture)."""
    import ctypes as C
    from move2kube_b200 import _abi, lib
    root = os.fsencode(treespec.materialize(spec, str(tmp_path / spec["name"])))
    fl = _abi.SH_BASH if flavor == "bash" else _abi.SH_DASH
    pre = [oracle.join(root, os.fsencode(p)) if p != "." else root for p in spec.get("pre", [])]
    tree = lib.pack_tree_host(root)
    try:
        ores = oracle.detect(tree, fl, pre)
        o, t = ores.contents, tree.contents
        view = gc.TreeView(tree)
        direct = (C.c_uint32 * max(1, t.n_dirs))()
        for e in range(t.n_entries):  # the one name feature m2k_option_json looks at: [ -f D/pom.xml ]
            if view.entry_name[e] == b"pom.xml" and (t.entry_type[e] >> 4) == _abi.T_REG:
                direct[t.entry_dir[e]] |= 1 << 3
        res = _abi.Result(n_dirs=t.n_dirs, n_files=t.n_files, n_entries=t.n_entries, dir_opts=o.dir_opts,
                          dir_direct=C.cast(direct, _abi.u32p), dir_main_file=o.dir_main_file)
        n = gc.check_outputs(spec, flavor, view, tree, C.pointer(res))
        assert n == sum(len(v) for v in spec["expect"][flavor]["outputs"].values())
        oracle.lib().m2ko_result_free(ores)
    finally:
        lib.tree_free_host(tree)
