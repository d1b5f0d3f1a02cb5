"""Host-only corners of the mirror (no GPU): what filepath.Walk does with a root that is missing or is
not a directory -- the first two cases of TestGetFilesByExt / TestGetFilesByName (utils_test.go:34-50, 90-107)."""
import os

from move2kube_b200 import common


def test_missing_root_is_an_error():
    files, err = common.GetFilesByExt("foobar-does-not-exist", [".yaml", ".yml"])
    assert err is not None
    files, err = common.GetFilesByName("foobar-does-not-exist", ["test1.yaml", "test2.yml"])
    assert err is not None


def test_root_that_is_a_file_is_reported_as_itself(tmp_path):
    d = tmp_path / "validfiles"
    d.mkdir()
    p = d / "test1.yaml"
    p.write_text("a: 1\n")
    assert common.GetFilesByExt(str(p), [".yaml", ".yml"]) == ([str(p)], None)
    assert common.GetFilesByName(str(p), ["test1.yaml", "test2.yml"]) == ([str(p)], None)
    assert common.GetFilesByExt(str(p), [".json"]) == ([], None)
    assert common.GetFilesByName(str(p), ["other"]) == ([], None)
    # Go's filepath.Ext: the suffix from the last dot of the base name, a leading dot included
    q = d / ".yaml"
    q.write_text("")
    assert common.GetFilesByExt(str(q), [".yaml"]) == ([str(q)], None)
    # a symlink to a directory is not descended into by Walk (Lstat): it is reported as itself
    link = tmp_path / "link.yml"
    os.symlink(str(d), str(link))
    assert common.GetFilesByExt(str(link), [".yml"]) == ([str(link)], None)
