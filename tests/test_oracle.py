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
