"""TEST INFRASTRUCTURE (oracle/) -- never imported by the product path.

Byte-level restatements of the Go standard-library helpers the reference's source-detection
path leans on.  [external]: the Go 1.15 stdlib is not under /root/reference; these follow its
documented behaviour and are pinned only by the reference fixtures that exercise them
(ASCII paths; see tests/test_oracle_gopath.py for the cases).

  clean / join / dirname / base   path/filepath (Unix)          -- call sites any2kube.go:165-176
  trim_space                      strings.TrimSpace             -- any2kube.go:168
  scan_lines                      bufio.Scanner + ScanLines     -- any2kube.go:164-167 (64 KiB token cap)
  equal_fold                      strings.EqualFold             -- internal/common/utils.go:308-315
"""
import unicodedata

MAX_SCAN_TOKEN = 64 * 1024  # bufio.MaxScanTokenSize


def clean(p: bytes) -> bytes:
    """path/filepath.Clean on Unix."""
    if p == b"":
        return b"."
    rooted = p.startswith(b"/")
    out = []
    for comp in p.split(b"/"):
        if comp == b"" or comp == b".":
            continue
        if comp == b"..":
            if out and out[-1] != b"..":
                out.pop()
            elif not rooted:
                out.append(b"..")
            continue
        out.append(comp)
    s = b"/".join(out)
    if rooted:
        return b"/" + s
    return s if s else b"."


def join(*elems: bytes) -> bytes:
    """path/filepath.Join: drops empty elements, then Clean; all-empty -> ''."""
    parts = [e for e in elems if e != b""]
    if not parts:
        return b""
    return clean(b"/".join(parts))


def dirname(p: bytes) -> bytes:
    """path/filepath.Dir."""
    i = p.rfind(b"/")
    d = clean(p[: i + 1])
    return d


def base(p: bytes) -> bytes:
    """path/filepath.Base."""
    if p == b"":
        return b"."
    while len(p) > 0 and p.endswith(b"/"):
        p = p[:-1]
    i = p.rfind(b"/")
    if i >= 0:
        p = p[i + 1:]
    if p == b"":
        return b"/"
    return p


_SPACE_RUNES = {0x09, 0x0A, 0x0B, 0x0C, 0x0D, 0x20, 0x85, 0xA0, 0x1680, 0x2028, 0x2029,
                0x202F, 0x205F, 0x3000} | set(range(0x2000, 0x200B))


def is_space_rune(r: int) -> bool:
    """unicode.IsSpace."""
    return r in _SPACE_RUNES


def decode_rune(b: bytes, i: int):
    """utf8.DecodeRune at b[i:]: returns (rune, width); invalid -> (0xFFFD, 1)."""
    n = len(b)
    c = b[i]
    if c < 0x80:
        return c, 1
    if 0xC2 <= c <= 0xDF:
        need, lo, hi, r = 1, 0x80, 0xBF, c & 0x1F
    elif c == 0xE0:
        need, lo, hi, r = 2, 0xA0, 0xBF, c & 0x0F
    elif 0xE1 <= c <= 0xEC or 0xEE <= c <= 0xEF:
        need, lo, hi, r = 2, 0x80, 0xBF, c & 0x0F
    elif c == 0xED:
        need, lo, hi, r = 2, 0x80, 0x9F, c & 0x0F
    elif c == 0xF0:
        need, lo, hi, r = 3, 0x90, 0xBF, c & 0x07
    elif 0xF1 <= c <= 0xF3:
        need, lo, hi, r = 3, 0x80, 0xBF, c & 0x07
    elif c == 0xF4:
        need, lo, hi, r = 3, 0x80, 0x8F, c & 0x07
    else:
        return 0xFFFD, 1
    if i + need >= n:  # truncated sequence
        return 0xFFFD, 1
    c1 = b[i + 1]
    if not (lo <= c1 <= hi):
        return 0xFFFD, 1
    r = (r << 6) | (c1 & 0x3F)
    for k in range(2, need + 1):
        ck = b[i + k]
        if not (0x80 <= ck <= 0xBF):
            return 0xFFFD, 1
        r = (r << 6) | (ck & 0x3F)
    return r, need + 1


def decode_last_rune(b: bytes, end: int):
    """utf8.DecodeLastRune on b[:end]: returns (rune, width)."""
    if end == 0:
        return 0xFFFD, 0
    start = end - 1
    c = b[start]
    if c < 0x80:
        return c, 1
    lim = max(0, end - 4)
    s = start - 1
    while s >= lim:
        if (b[s] & 0xC0) != 0x80:  # rune start
            break
        s -= 1
    if s < lim:
        s = lim
    r, w = decode_rune(b[:end], s)
    if s + w != end:
        return 0xFFFD, 1
    return r, w


def trim_space(s: bytes) -> bytes:
    """strings.TrimSpace."""
    i, n = 0, len(s)
    while i < n:
        r, w = decode_rune(s, i)
        if not is_space_rune(r):
            break
        i += w
    j = n
    while j > i:
        r, w = decode_last_rune(s[i:], j - i)
        if not is_space_rune(r):
            break
        j -= w
    return s[i:j]


def scan_lines(data: bytes):
    """bufio.Scanner with ScanLines and the default 64 KiB buffer cap.

    Yields each line without its '\\n' and without one trailing '\\r'.  A line whose content
    (including a trailing '\\r') is >= 65536 bytes makes the scanner stop with ErrTooLong:
    nothing further is yielded.  Returns via StopIteration value True if ErrTooLong occurred.
    """
    pos, n = 0, len(data)
    while pos < n:
        nl = data.find(b"\n", pos)
        end = n if nl < 0 else nl
        if end - pos >= MAX_SCAN_TOKEN:
            return True
        line = data[pos:end]
        if line.endswith(b"\r"):
            line = line[:-1]
        yield line
        pos = end + 1
    return False


# ---- strings.EqualFold -------------------------------------------------------------------

# Case pairs that entered Unicode after 12.0 (Go 1.15 ships Unicode 12.0 tables): leave unfolded.
_POST_U12 = (set(range(0xA7C7, 0xA7CB)) | {0xA7F5, 0xA7F6, 0x2C2F, 0x2C5F, 0xA7C0, 0xA7C1,
             0xA7D0, 0xA7D1} | set(range(0xA7D6, 0xA7DA)) | set(range(0x10570, 0x105BD)))


def _simple_fold_key(cp: int) -> int:
    """Representative of cp's simple-case-folding class (CaseFolding.txt statuses C+S).

    Go's unicode.SimpleFold walks exactly these classes (caseOrbit + ToLower/ToUpper, with
    U+0130 / U+0131 pinned to themselves).  Python only exposes *full* folding, so where the
    full folding is multi-character we fall back to the single-character lowercase mapping.
    """
    if cp < 0x80:
        return cp + 32 if 0x41 <= cp <= 0x5A else cp
    if cp in _POST_U12 or 0xD800 <= cp <= 0xDFFF:
        return cp
    ch = chr(cp)
    cf = ch.casefold()
    if len(cf) == 1:
        return ord(cf)
    lo = ch.lower()
    if len(lo) == 1:
        return ord(lo)
    return cp


_fold_cache = {}


def fold_key(cp: int) -> int:
    k = _fold_cache.get(cp)
    if k is None:
        k = _simple_fold_key(cp)
        if k != cp:
            # the key itself must be a fixed point so that classes are well defined
            k2 = _simple_fold_key(k)
            if k2 != k:
                k = k2
        _fold_cache[cp] = k
    return k


def fold_runes(s: bytes):
    """Canonical folded rune sequence: EqualFold(a, b) <=> fold_runes(a) == fold_runes(b)."""
    out = []
    i, n = 0, len(s)
    while i < n:
        r, w = decode_rune(s, i)
        out.append(fold_key(r))
        i += w
    return out


def equal_fold(a: bytes, b: bytes) -> bool:
    return fold_runes(a) == fold_runes(b)


def fold_table():
    """[(codepoint, key)] for every non-ASCII codepoint whose key differs from itself."""
    out = []
    for cp in range(0x80, 0x110000):
        k = fold_key(cp)
        if k != cp:
            out.append((cp, k))
    return out


UNIDATA_VERSION = unicodedata.unidata_version
