"""The device-side isDockerFile source, compiled for the host, against the oracle -- by the hundred thousand.

`tests/host_df/df_host.cc` includes `move2kube_b200/csrc/dockerfile_dev.cuh` (the file nvcc compiles into k_df_head /
k_df_states / k_df_validate) with `__device__` defined away and drives it the way the three kernels do.  The GPU tier
runs a few thousand Dockerfile-shaped inputs per box-minute; here the same parsing code meets a grammar-driven fuzz of
several hundred thousand, at stripe widths from one byte (every line start is a stripe boundary) to the shipped one.
The oracle (oracle/m2k_oracle.c m2ko_dockerfile_flags) is the other restatement of
internal/source/dockerfile2kube.go:117-144 + buildkit v0.7.2's parser; a disagreement is a bug in one of them.
Nothing here is on the product path: the product runs the CUDA build of the same header.
"""
import ctypes as C
import os
import random
import subprocess

import pytest

import oracle_lib
from dockerfile_kats import KATS, reference_corpus

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
F_LONG, F_FROM, F_DF = 2, 4, 8
MASK = F_LONG | F_FROM | F_DF


@pytest.fixture(scope="module")
def dfhost():
    out_dir = os.path.join(HERE, "host_df", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libdfhost.so")
    srcs = [os.path.join(HERE, "host_df", "df_host.cc"), os.path.join(ROOT, "move2kube_b200", "csrc", "dockerfile_dev.cuh"),
            os.path.join(ROOT, "move2kube_b200", "csrc", "gostr.h"), os.path.join(ROOT, "move2kube_b200", "csrc", "runacc.h")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-Wall", "-Wno-unknown-pragmas", "-Wno-sign-compare",
                               "-I", os.path.join(ROOT, "move2kube_b200", "csrc"), "-I", os.path.join(ROOT, "include"), srcs[0], "-o", so])
    L = C.CDLL(so)
    L.dfhost_flags.restype = C.c_uint32
    L.dfhost_flags.argtypes = [C.c_char_p, C.c_uint64, C.c_uint32]
    L.dfhost_head_prefix.restype = C.c_uint32
    L.dfhost_head_prefix.argtypes = [C.c_char_p, C.c_uint64]
    L.dfhost_default_stripe.restype = C.c_uint32
    L.dfhost_runacc_check.restype = C.c_uint32
    L.dfhost_runacc_check.argtypes = [C.c_char_p, C.c_uint64, C.POINTER(C.c_uint64), C.c_uint32, C.c_uint64, C.c_uint32]
    return L


# ---- the fuzz grammar ---------------------------------------------------------------------------------------------
WS = [b" ", b" ", b" ", b"\t", b"  ", b"\x0b", b"\x0c", b"\r", b"\xc2\xa0", b"\xc2\x85", b"\xe2\x80\x83", b"\x85", b"\xa0"]
KEYWORDS = [b"FROM", b"from", b"From", b"ARG", b"arg", b"ENV", b"env", b"LABEL", b"label", b"ONBUILD", b"onbuild", b"RUN", b"CMD",
            b"COPY", b"HEALTHCHECK", b"FROMX", b"EN", b"V", b"FR", b"OM", b"ONBU\xc4\xb0LD", b"\xe2\x84\xaa", b"LABE", b"L"]
WORDS = [b"a", b"b", b"ubuntu:20.04", b"x=y", b"A=1", b"B", b"=", b"==", b"k=\"v w\"", b"'q r'=s", b"\"", b"'", b"--flag", b"--platform=linux/amd64",
         b"--", b"-", b"AS", b"as", b"A\xc5\xbf", b"build", b"#", b"#c", b"# trailing", b"\\", b"`", b"\\ ", b"` ", b"\\\\", b"$V", b"[\"x\"]",
         b"\xff", b"\xc3", b"\xe2\x82", b"\x00", b"caf\xc3\xa9", b"a#b", b"x\\", b"y`"]
EOL = [b"\n", b"\n", b"\n", b"\r\n", b"\\\n", b"\\ \t\n", b"`\n", b"\\\r\n", b"\n\n", b" \n", b"\\\n#in\n", b"\\\n\n", b"\\\n \t\n"]
DIRECTIVES = [b"# escape=`\n", b"#escape=\\\n", b"# escape = `\n", b"# ESCAPE=`\r\n", b"# escape=x\n", b"# escape= \n", b"# escape=\n", b"# syntax=docker/dockerfile:1\n",
              b"# hello\n", b"#\n", b"\xef\xbb\xbf", b"\xef\xbb\xbf# escape=`\n", b"   # escape=`\n"]


def gen_line(r):
    parts = []
    if r.random() < 0.25:
        parts.append(r.choice(WS))
    kind = r.random()
    if kind < 0.08:
        parts.append(b"#" + r.choice(WORDS))
    elif kind < 0.14:
        pass  # blank line
    else:
        parts.append(r.choice(KEYWORDS))
        for _ in range(r.randrange(0, 6)):
            if r.random() < 0.9:
                parts.append(r.choice(WS))
            parts.append(r.choice(WORDS) if r.random() < 0.85 else r.choice(KEYWORDS))
    if r.random() < 0.2:
        parts.append(r.choice(WS))
    parts.append(r.choice(EOL))
    return b"".join(parts)


def gen_file(r):
    mode = r.random()
    out = []
    if mode < 0.15:  # token soup
        pool = WS + KEYWORDS + WORDS + EOL
        return b"".join(r.choice(pool) for _ in range(r.randrange(0, 40)))
    for _ in range(r.randrange(0, 3)):
        if r.random() < 0.5:
            out.append(r.choice(DIRECTIVES))
    for _ in range(r.randrange(0, 3)):
        if r.random() < 0.5:
            out.append(b"ARG " + r.choice(WORDS) + r.choice(EOL))
    if r.random() < 0.85:  # a FROM line that mostly matches, so that the body gets looked at
        out.append(r.choice([b"FROM ", b"from\t", b"FROM  ", b"FR\\\nOM ", b"FROM \\\n "]) + r.choice([b"a", b"ubuntu:20.04", b"$V", b"a#b"]) +
                   r.choice([b"", b"", b" AS b", b" as  x", b" A\xc5\xbf y", b" # c", b" b", b" AS"]) + r.choice(EOL))
    for _ in range(r.randrange(0, 12)):
        out.append(gen_line(r))
    data = b"".join(out)
    if r.random() < 0.3 and data.endswith(b"\n"):
        data = data[:-1]  # no newline at the end
    return data


def check(dfhost, data, stripe):
    want = oracle_lib.dockerfile_flags(data) & MASK
    got = dfhost.dfhost_flags(data, len(data), stripe)
    assert not (got >> 31), ("map composition is not associative on", data)
    assert got == want, (data, stripe, "host build of the device code", got, "oracle", want)


def test_flag_values_match_the_header():
    from move2kube_b200 import _abi
    assert (_abi.FILE_LONG_LINE, _abi.FILE_FROM_FIRST, _abi.FILE_DOCKERFILE) == (F_LONG, F_FROM, F_DF)


def test_kats_and_reference_corpus_through_the_host_build(dfhost):
    assert dfhost.dfhost_default_stripe() == 128
    for name, data, want in KATS:
        for stripe in (0, 1, 5, 64):
            got = dfhost.dfhost_flags(data, len(data), stripe)
            assert bool(got & F_DF) == want, (name, stripe)
    for name, data, want in reference_corpus():
        for stripe in (0, 3):
            check(dfhost, data, stripe)
            assert bool(dfhost.dfhost_flags(data, len(data), stripe) & F_DF) == want, name


@pytest.mark.parametrize("seed", range(8))
def test_fuzz_against_the_oracle(dfhost, seed):
    r = random.Random(0xD0C0 + seed)
    stripes = (1, 2, 3, 7, 16, 0)
    n_df = 0
    for i in range(30000):
        data = gen_file(r)
        check(dfhost, data, stripes[i % len(stripes)])
        n_df += bool(oracle_lib.dockerfile_flags(data) & F_DF)
    assert 1500 < n_df < 28000  # the grammar reaches both answers often enough to mean something


def test_first_words(dfhost):
    """What the byte scan's pre-filter sees -- the first 16 bytes -- in every pairing of the grammar's tokens, and files
    that start with any line the grammar makes (no FROM scaffold in front)."""
    pool = WS + KEYWORDS + WORDS + EOL + [b"\xef\xbb\xbf", b"\n\n\n\n\n\n\n\n", b" " * 15, b" " * 16, b"\t" * 12]
    n = 0
    for a in pool:
        for b in pool:
            for tail in (b" a\n", b"FROM a\n", b""):
                check(dfhost, a + b + tail, (0, 2)[n & 1])
                n += 1
    r = random.Random(31)
    for _ in range(60000):
        check(dfhost, b"".join(gen_line(r) for _ in range(r.randrange(1, 5))), 0)


def test_mutations_of_the_kats(dfhost):
    """Every known-answer vector with one byte dropped, doubled or replaced, at every position (short vectors)."""
    r = random.Random(7)
    repl = [b"\n", b"\\", b"`", b" ", b"#", b"=", b"\"", b"'", b"\r", b"\xc2\xa0", b"-", b"E"]
    n = 0
    for name, data, _ in KATS:
        if len(data) > 200:
            continue
        for pos in range(len(data) + 1):
            for mut in (data[:pos] + data[pos + 1:], data[:pos] + data[pos:pos + 1] + data[pos:], data[:pos] + r.choice(repl) + data[pos + 1:],
                        data[:pos] + r.choice(repl) + data[pos:]):
                check(dfhost, mut, (1, 4, 0)[n % 3])
                n += 1
    assert n > 5000


@pytest.mark.parametrize("seed", range(4))
def test_a_head_settled_on_a_prefix_stays_settled(dfhost, seed):
    """A GPU that holds only the first part of a cut file settles FROM-first from that part unless the verdict rests
    on the last line present (k_df_head, PIECE_HEAD_*): whatever the prefix settles must be what the whole file says."""
    r = random.Random(99 + seed)
    settled = 0
    for _ in range(20000):
        data = gen_file(r)
        if len(data) < 2 or (oracle_lib.dockerfile_flags(data) & F_LONG):
            continue
        whole = bool(oracle_lib.dockerfile_flags(data) & F_FROM)
        for cut in {r.randrange(1, len(data)) for _ in range(4)}:
            v = dfhost.dfhost_head_prefix(data[:cut], cut)
            if v == 0:
                continue
            settled += 1
            assert (v == 1) == whole, (data, cut, v)
    assert settled > 10000


def test_newline_runs_fold_the_same_in_any_grouping(dfhost):
    """LONG_LINE comes from per-chunk (lead, trail) records merged per thread, per block share of a big file and per
    piece of a file cut between GPUs (runacc.h): any grouping of any cutting of a span must give the summary of the
    whole, and the flag must say "some line has >= limit bytes".  Short limits, so that every boundary case is common."""
    r = random.Random(5)
    for case in range(40000):
        limit = r.choice((1, 2, 3, 5, 8, 13))
        n = r.randrange(0, 48)
        p_nl = r.choice((0.0, 0.05, 0.2, 0.5, 0.9))
        data = bytes(10 if r.random() < p_nl else 120 for _ in range(n))
        cuts = sorted({0, n} | {r.randrange(0, n + 1) for _ in range(r.randrange(0, 9))}) if n else [0]
        if r.random() < 0.2 and n:  # chunk-like: equal spans
            step = r.randrange(1, 9)
            cuts = list(range(0, n, step)) + [n]
        arr = (C.c_uint64 * len(cuts))(*cuts)
        rc = dfhost.dfhost_runacc_check(data, n, arr, len(cuts) - 1, limit, case)
        assert rc == 0, (rc, data, cuts, limit)
