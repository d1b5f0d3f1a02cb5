"""CPU tests: Go-stdlib restatements and the Dockerfile known answers, on the oracle."""
import pytest

import gopath
from dockerfile_kats import KATS, reference_corpus
from move2kube_b200 import _abi

CLEAN = [(b"", b"."), (b"/", b"/"), (b"a//b", b"a/b"), (b"a/./b/", b"a/b"), (b"a/b/..", b"a"), (b"../a", b"../a"),
         (b"a/../..", b".."), (b"/../a", b"/a"), (b"./", b"."), (b"a/b/../../..", b".."), (b"/a/b/../../..", b"/"),
         (b"abc/def/../ghi", b"abc/ghi"), (b"../../a/..", b"../.."), (b"a/.hidden/..", b"a")]


@pytest.mark.parametrize("src,want", CLEAN)
def test_clean(oracle, src, want):
    assert gopath.clean(src) == want
    assert oracle.clean(src) == want


def test_join(oracle):
    for a, b, want in [(b"/r/x", b"", b"/r/x"), (b"/r/x", b".", b"/r/x"), (b"/r/x", b"/excludeme/", b"/r/x/excludeme"),
                       (b"/r/x", b"../y", b"/r/y"), (b"/r", b"a/*", b"/r/a/*"), (b"", b"", b"")]:
        assert gopath.join(a, b) == want
        assert oracle.join(a, b) == want


FOLD = [(b"Go", b"GO", True), (b"K", "K".encode(), True), (b"s", "ſ".encode(), True), (b"I", "ı".encode(), False),
        (b"i", "İ".encode(), False), ("µ".encode(), "μ".encode(), True), ("ß".encode(), "ẞ".encode(), True),
        (b"a\xffb", b"A\xfeB", True), (b"\xef\xbf\xbd", b"\x80", True), (b"abc", b"abcd", False), (b"/a/b", b"/A/B", True),
        ("ǅ".encode(), "ǆ".encode(), True), (b"ss", "ß".encode(), False)]


@pytest.mark.parametrize("a,b,want", FOLD)
def test_equal_fold(oracle, a, b, want):
    assert gopath.equal_fold(a, b) == want
    assert oracle.equal_fold(a, b) == want
    assert oracle.equal_fold(b, a) == want


def test_trim_space_and_scan_lines():
    assert gopath.trim_space(b" \t a b \r\n") == b"a b"
    assert gopath.trim_space(" x ".encode()) == b"x"
    assert gopath.trim_space(b"\x85x") == b"\x85x"  # a lone 0x85 byte is not U+0085
    assert list(gopath.scan_lines(b"a\r\nb\n\nc")) == [b"a", b"b", b"", b"c"]
    assert list(gopath.scan_lines(b"a\n" + b"x" * 65536 + b"\nb\n")) == [b"a"]
    assert list(gopath.scan_lines(b"x" * 65535 + b"\nb")) == [b"x" * 65535, b"b"]


@pytest.mark.parametrize("name,data,want", KATS, ids=[k[0] for k in KATS])
def test_dockerfile_kat(oracle, name, data, want):
    flags = oracle.file_flags(data)
    assert bool(flags & _abi.FILE_DOCKERFILE) == want, "flags=%#x" % flags
    assert bool(flags & _abi.FILE_HAS_MAIN) == (b"__main__" in data)


CORPUS = reference_corpus()


@pytest.mark.parametrize("path,data,want", CORPUS, ids=[c[0] for c in CORPUS])
def test_dockerfiles_that_ship_in_the_reference_tree(oracle, path, data, want):
    """Every real Dockerfile of the reference tree must be accepted, a spread of its other files rejected
    (expectations by inspection: tests/golden/make_dockerfile_corpus.py)."""
    flags = oracle.file_flags(data)
    assert bool(flags & _abi.FILE_DOCKERFILE) == want, "%s flags=%#x" % (path, flags)
