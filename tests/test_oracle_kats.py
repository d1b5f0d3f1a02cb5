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


def test_fold_table_agrees_with_a_second_derivation(oracle):
    """strings.EqualFold beyond ASCII rests on one generated table (tools/gen_casefold.py, from str.casefold()), which
    the oracle and the product both compile in.  This derives the classes a SECOND way -- the closure of the Unicode
    database's single-character upper / lower / title mappings, which is what Go's unicode.SimpleFold walks -- and
    checks both compiled tables against it for every cased code point.  The only differences allowed are the ones
    Go itself makes: dotless / dotted i (U+0131, U+0130) fold to themselves, and pairs that entered Unicode after
    12.0 (Go 1.15's tables) are not folded."""
    from move2kube_b200 import lib
    parent = {}

    def find(x):
        while parent.get(x, x) != x:
            parent[x] = parent.get(parent[x], parent[x])
            x = parent[x]
        return x

    for cp in range(0x41, 0x110000):
        if 0xD800 <= cp <= 0xDFFF:
            continue
        ch = chr(cp)
        for m in (ch.upper(), ch.lower(), ch.title()):
            if len(m) == 1 and ord(m) != cp:
                ra, rb = find(cp), find(ord(m))
                if ra != rb:
                    parent[max(ra, rb)] = min(ra, rb)
    groups = {}
    for cp in set(parent) | set(parent.values()):
        groups.setdefault(find(cp), set()).add(cp)
    post_u12 = gopath._POST_U12
    checked = 0
    for g in groups.values():
        members = sorted(g)
        for a in members[1:]:
            b = members[0]
            sa, sb = chr(a).encode("utf-8"), chr(b).encode("utf-8")
            want = not ({a, b} & {0x130, 0x131}) and not ({a, b} & post_u12)
            assert oracle.equal_fold(sa, sb) == want, "oracle: U+%04X vs U+%04X" % (a, b)
            assert lib.equal_fold(sa, sb) == want, "product: U+%04X vs U+%04X" % (a, b)
            checked += 1
    assert checked > 1300
    # and nothing outside those classes is folded together: neighbours in code point order that the closure keeps apart
    for a in range(0x80, 0x3000):
        if find(a) != find(a + 1):
            assert not lib.equal_fold(chr(a).encode("utf-8"), chr(a + 1).encode("utf-8")), "U+%04X / U+%04X" % (a, a + 1)
