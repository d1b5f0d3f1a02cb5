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


def _build_dfhost(name, extra=()):
    out_dir = os.path.join(HERE, "host_df", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, name)
    srcs = [os.path.join(HERE, "host_df", "df_host.cc"), os.path.join(ROOT, "move2kube_b200", "csrc", "dockerfile_dev.cuh"),
            os.path.join(ROOT, "move2kube_b200", "csrc", "gostr.h"), os.path.join(ROOT, "move2kube_b200", "csrc", "runacc.h")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-Wall", "-Wno-unknown-pragmas", "-Wno-sign-compare", *extra,
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


@pytest.fixture(scope="module")
def dfhost():
    return _build_dfhost("libdfhost.so")


@pytest.fixture(scope="module")
def dfhost_keep2():
    """the same source with only TWO levels of JSON nesting kept in the register (the product keeps 64): the pass that
    recovers the kinds of the enclosing containers then runs all the time instead of practically never"""
    return _build_dfhost("libdfhost_keep2.so", ("-DM2K_JSON_KEEP=2",))


# ---- the fuzz grammar ---------------------------------------------------------------------------------------------
WS = [b" ", b" ", b" ", b"\t", b"  ", b"\x0b", b"\x0c", b"\r", b"\xc2\xa0", b"\xc2\x85", b"\xe2\x80\x83", b"\x85", b"\xa0"]
KEYWORDS = [b"FROM", b"from", b"From", b"ARG", b"arg", b"ENV", b"env", b"LABEL", b"label", b"ONBUILD", b"onbuild", b"RUN", b"CMD",
            b"COPY", b"HEALTHCHECK", b"FROMX", b"EN", b"V", b"FR", b"OM", b"ONBU\xc4\xb0LD", b"\xe2\x84\xaa", b"LABE", b"L"]
WORDS = [b"a", b"b", b"ubuntu:20.04", b"x=y", b"A=1", b"B", b"=", b"==", b"k=\"v w\"", b"'q r'=s", b"\"", b"'", b"--flag", b"--platform=linux/amd64",
         b"--", b"-", b"AS", b"as", b"A\xc5\xbf", b"build", b"#", b"#c", b"# trailing", b"\\", b"`", b"\\ ", b"` ", b"\\\\", b"$V", b"[\"x\"]",
         b"\xff", b"\xc3", b"\xe2\x82", b"\x00", b"caf\xc3\xa9", b"a#b", b"x\\", b"y`"]
EOL = [b"\n", b"\n", b"\n", b"\r\n", b"\\\n", b"\\ \t\n", b"`\n", b"\\\r\n", b"\n\n", b" \n", b"\\\n#in\n", b"\\\n\n", b"\\\n \t\n",
       b"\\\n  # indented\n", b"\\\n\t#x \\\n", b"\\\n\xc2\xa0#y\n", b"`\n  # in\n"]
DIRECTIVES = [b"# escape=`\n", b"#escape=\\\n", b"# escape = `\n", b"# ESCAPE=`\r\n", b"# escape=x\n", b"# escape= \n", b"# escape=\n", b"# syntax=docker/dockerfile:1\n",
              b"# hello\n", b"#\n", b"\xef\xbb\xbf", b"\xef\xbb\xbf# escape=`\n", b"   # escape=`\n"]


# arguments in JSON form (parseMaybeJSON*): well formed and not, strings only and not, numbers in and out of float64 range
JSON_CMDS = [b"RUN", b"CMD", b"run", b"COPY", b"ADD", b"ENTRYPOINT", b"SHELL", b"VOLUME", b"HEALTHCHECK CMD", b"HEALTHCHECK --interval=5s CMD",
             b"HEALTHCHECK NONE", b"ONBUILD RUN", b"ONBUILD ONBUILD COPY", b"EXPOSE", b"ENV", b"Cmd", b"ENTRYPO\xc4\xb0NT", b"HEALTHCHEC\xe2\x84\xaa X",
             b"RUN --mount=type=cache", b"COPY --from=0 --chown=1:1"]
JSON_ATOMS = [b'"a"', b'"b c"', b'""', b'"\\""', b'"\\u00e9"', b'"\\u12G4"', b'"\\x"', b'"tab\there"', b'"caf\xc3\xa9"', b'"\xff"', b"1", b"0", b"-0", b"01", b"-", b"1.5",
              b"1.", b".5", b"1e5", b"1E+5", b"1e", b"1e999", b"-1e400", b"1.7976931348623157e308", b"1.7976931348623158e308", b"1.7976931348623159e308",
              b"17976931348623158" + b"0" * 292, b"0." + b"0" * 30 + b"1e340", b"1e-999", b"true", b"false", b"null", b"tru", b"nul", b"True", b"[]", b"[ ]", b'["x"]',
              b"[1]", b"{}", b'{"k":1}', b'{"k":"v"}', b'{"k"}', b'{"k":}', b"{1:2}", b'{"a":[{"b":[null]}]}', b"'a'", b"a", b"[[[[[[[[[[]]]]]]]]]]"]
JSON_SEPS = [b",", b",", b",", b", ", b" ,", b" , ", b"", b" ", b",,", b":", b",\t", b",\\\n ", b",\\\n  # c\n "]


def gen_json_line(r):
    cmd = r.choice(JSON_CMDS)
    sep = r.choice([b" ", b" ", b"\t", b"  ", b" \xc2\xa0", b" \\\n "])
    n = r.choice((0, 1, 1, 2, 2, 3, 5))
    body = b""
    for k in range(n):
        if k:
            body += r.choice(JSON_SEPS)
        body += r.choice(JSON_ATOMS)
    arr = r.choice([b"[", b"[", b"[", b"[ ", b" [", b"[[", b""]) + body + r.choice([b"]", b"]", b"]", b" ]", b"] tail", b"]]", b"", b"]#c", b",]"])
    return cmd + sep + arr + r.choice([b"\n", b"\n", b" \n", b"\r\n", b""])


def deep_json(depth, kinds, leaf=b"1", close=True):
    """nested containers: kinds[i % len] picks '[' or '{"k":' per level"""
    opens, closes = [], []
    for i in range(depth):
        if kinds[i % len(kinds)] == "a":
            opens.append(b"[")
            closes.append(b"]")
        else:
            opens.append(b'{"k":')
            closes.append(b"}")
    return b"".join(opens) + leaf + (b"".join(reversed(closes)) if close else b"")


def gen_line(r):
    if r.random() < 0.18:
        return gen_json_line(r)
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


def _third():
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import dockerfile_re
    return dockerfile_re


_checked = [0]


def check(dfhost, data, stripe):
    want = oracle_lib.dockerfile_flags(data) & MASK
    got = dfhost.dfhost_flags(data, len(data), stripe)
    assert not (got >> 31), ("map composition is not associative on", data)
    assert got == want, (data, stripe, "host build of the device code", got, "oracle", want)
    _checked[0] += 1
    if _checked[0] % 3 == 0:  # every third case also through the regex-engine derivation (it is the slow one)
        assert _third().is_dockerfile(data) == bool(want & F_DF), (data, "regex-engine derivation vs oracle", want)


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


# ---- arguments in JSON form: three derivations ----------------------------------------------------------------------
from dockerfile_kats import JSON_KATS  # noqa: E402


def test_json_kats_three_ways(dfhost, dfhost_keep2):
    third = _third()
    for data, want in JSON_KATS:
        assert third.is_dockerfile(data) == want, ("regex-engine derivation", data)
        assert bool(oracle_lib.dockerfile_flags(data) & F_DF) == want, ("C oracle", data)
        for lib in (dfhost, dfhost_keep2):
            for stripe in (0, 1, 7):
                assert bool(lib.dfhost_flags(data, len(data), stripe) & F_DF) == want, ("host build of the device code", data, stripe)


def test_float64_limit_digit_by_digit(dfhost):
    """the device compares a number literal with 2^1024 - 2^970 digit by digit; Python's float() rounds correctly"""
    limit = 2 ** 1024 - 2 ** 970
    r = random.Random(11)
    lits = [str(limit), str(limit - 1), str(limit + 1), str(limit)[:-1] + "1.999", str(limit)[:200] + "e109", str(limit)[:200] + "0e108",
            "0." + str(limit) + "e309", "0.0" + str(limit) + "e310", "0.00" + str(limit - 1) + "e311", str(limit) + "0e-1", str(limit - 1) + "00E-2",
            "1" + "0" * 308, "1" + "0" * 309, "9" * 308, "9" * 309, "1e308", "2e308", "0e999", "0.0e999", "-0.000e+99999999999999999999", "1e99999999999999999999",
            "1e-99999999999999999999", "17976931348623158e292", "17976931348623159e292", "0.00017976931348623159e312"]
    for _ in range(4000):
        # the limit plus or minus something, cut to a random number of digits, the point put anywhere, the exponent making up for it
        digits = str(limit + r.choice((-1, 0, 0, 1)) * r.randrange(0, 10 ** r.randrange(0, 300)))
        cut = r.randrange(1, len(digits) + 1)
        frac_at = r.randrange(0, cut + 1)
        exp = len(digits) - frac_at + r.choice((0, 0, 0, 0, -1, 1))
        zeros = "0" * r.choice((0, 0, 1, 5)) if frac_at == 0 else ""   # 0.000ddd: zeros after the point, made up for in the exponent
        lit = (digits[:frac_at] or "0") + ("." + zeros + digits[frac_at:cut] if frac_at < cut else "") + r.choice("eE") + r.choice(("", "+")) + str(exp + len(zeros))
        lits.append(lit)
    n_inf = 0
    for lit in lits:
        if lit[0] == "0" and len(lit) > 1 and lit[1].isdigit():
            continue  # not a JSON number
        data = b"FROM a\nRUN [" + lit.encode() + b"]\n"
        inf = float(lit) in (float("inf"), float("-inf"))
        n_inf += inf
        got = bool(dfhost.dfhost_flags(data, len(data), 0) & F_DF)
        assert got == inf, (lit, "out of range -> Decode fails -> plain string -> still a Dockerfile" if inf else "in range -> not a string -> Parse fails")
        assert bool(oracle_lib.dockerfile_flags(data) & F_DF) == inf, ("C oracle", lit)
    assert 300 < n_inf < len(lits) - 300


@pytest.mark.parametrize("seed", range(3))
def test_json_lines_fuzz_three_ways(dfhost, dfhost_keep2, seed):
    third = _third()
    r = random.Random(500 + seed)
    n_bad = 0
    for i in range(25000):
        data = b"FROM a\n" + b"".join(gen_json_line(r) if r.random() < 0.8 else gen_line(r) for _ in range(r.randrange(1, 4)))
        want = third.is_dockerfile(data)
        n_bad += not want
        assert bool(oracle_lib.dockerfile_flags(data) & F_DF) == want, ("C oracle vs regex-engine derivation", data)
        lib = (dfhost, dfhost_keep2)[i & 1]
        got = lib.dfhost_flags(data, len(data), (0, 1, 5)[i % 3])
        assert bool(got & F_DF) == want, ("host build of the device code", data)
    assert 2000 < n_bad < 23000


def test_json_nesting_depth(dfhost, dfhost_keep2):
    """Go's scanner gives up beyond 10000 open containers (Decode fails: plain string); the device keeps the kinds of the
    innermost 64 (here also: 2) levels in a register and recovers the others by a second pass when it climbs back out"""
    third = _third()
    cases = []
    for depth, kinds in ((1, "a"), (2, "o"), (3, "ao"), (5, "oa"), (63, "a"), (64, "ao"), (65, "oao"), (66, "o"), (130, "aao"), (300, "aoo"), (1000, "oa"),
                         (9999, "a"), (10000, "a"), (10001, "a"), (10000, "ao"), (10001, "oa")):
        for leaf in (b"1", b'"s"', b"[]", b"{}"):
            inner = deep_json(depth - 1, kinds, leaf)
            cases.append(b"FROM a\nRUN [" + inner + b"]\n")
            cases.append(b"FROM a\nRUN [" + inner + b', "x", ' + inner + b"]\n")
            cases.append(b"FROM a\nRUN [" + inner[:-1] + b"]\n" if depth > 1 else b"FROM a\nRUN [\n")   # one closer short, or mismatched
            cases.append(b"FROM a\nRUN [" + inner.replace(b"}", b"]", 1) + b"]\n")                  # a wrong kind of closer somewhere
    r = random.Random(3)
    for _ in range(400):  # random walks up and down
        out, stack = [b"["], ["a"]
        for _ in range(r.randrange(1, 400)):
            top = stack[-1] if stack else None
            if not stack:
                break
            act = r.random()
            prev = out[-1]
            need_sep = prev not in (b"[", b"{", b",", b":") and not prev.endswith(b":")
            if act < 0.45 and len(stack) < 200:
                if need_sep:
                    out.append(b",")
                if top == "o":
                    out.append(b'"k":')
                k = r.choice("ao")
                out.append(b"[" if k == "a" else b"{")
                stack.append(k)
            elif act < 0.75:
                if need_sep:
                    out.append(b",")
                if top == "o":
                    out.append(b'"k":')
                out.append(r.choice([b"1", b'"s"', b"null", b'"]"', b'"\\"}"']))
            else:
                if prev in (b",",):
                    out.pop()
                closer = b"]" if top == "a" else b"}"
                if r.random() < 0.03:
                    closer = b"}" if top == "a" else b"]"
                out.append(closer)
                stack.pop()
        while stack and r.random() < 0.9:
            if out[-1] == b",":
                out.pop()
            out.append(b"]" if stack.pop() == "a" else b"}")
        cases.append(b"FROM a\nRUN " + b"".join(out) + b"\n")
    n_bad = 0
    for data in cases:
        want = third.is_dockerfile(data)
        n_bad += not want
        assert bool(oracle_lib.dockerfile_flags(data) & F_DF) == want, ("C oracle", data[:120], len(data))
        for lib in (dfhost, dfhost_keep2):
            assert bool(lib.dfhost_flags(data, len(data), 0) & F_DF) == want, ("host build of the device code", data[:120], len(data))
    assert 50 < n_bad < len(cases) - 50
