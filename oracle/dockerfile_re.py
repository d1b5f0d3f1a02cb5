"""TEST INFRASTRUCTURE -- a THIRD derivation of isDockerFile (internal/source/dockerfile2kube.go:117-144), kept as
literal as Python allows.

The other two (oracle/m2k_oracle.c m2ko_dockerfile_flags and move2kube_b200/csrc/dockerfile_dev.cuh) are hand-written
state machines; this one lets a regular-expression ENGINE do what the Go code does with regular expressions, builds the
strings the Go code builds (the joined line, the words, the flags) and asks the questions the Go code asks of them
(`strings.Contains(word, "=")`, `len(parts) < 2`).  Same author, different method: a slip in a hand-rolled loop (how a
greedy `[ \\t]*` hands a blank back to `.`, what `$` sees after trimming, where a run of non-space ends) shows up as a
disagreement.  It is slow and only used by the CPU tier of the tests.

[external] moby/buildkit v0.7.2 (go.mod:22) frontend/dockerfile/parser: parser.go (Parse, processLine,
possibleParserDirective, trimContinuationCharacter, isComment, isEmptyContinuationLine, newNodeFromLine), split_command.go
(splitCommand, extractBuilderFlags), line_parsers.go (parseWords, parseNameVal, parseSubCommand, parseIgnore).  The source
is not in the reference tree; this follows the published algorithm, function by function.  Parity: unpinned.

Go strings are bytes: everything here works on `bytes`, decoding runes where Go does (invalid UTF-8 = U+FFFD, width 1).
"""
import re

MAX_SCAN_TOKEN = 65536  # bufio.MaxScanTokenSize

RE_S = rb"[\t\n\f\r ]"          # RE2 \s
RE_NS = rb"[^\t\n\f\r ]"        # [^\s]
TOKEN_WHITESPACE = re.compile(rb"[\t\v\f\r ]+")
# ^#[ \t]*escape[ \t]*=[ \t]*(?P<escapechar>.).*$   (applied to strings.ToLower(line); a line holds no '\n')
TOKEN_ESCAPE_COMMAND = re.compile(rb"\A#[ \t]*escape[ \t]*=[ \t]*(?P<escapechar>(?:[\x00-\x7f]|[\xc2-\xf4][\x80-\xbf]*|[\x80-\xff]))[^\n]*\Z")
TOKEN_COMMENT = re.compile(rb"\A#[^\n]*\Z")
# the reference's own pattern, dockerfile2kube.go:131.  (?i) folds s with U+017F (and k with U+212A, not in the pattern)
FROM_RE = re.compile(rb"(?:[Ff][Rr][Oo][Mm])" + RE_S + rb"+(?:--[Pp][Ll][Aa][Tt][Ff][Oo][Rr][Mm]=" + RE_NS + rb"+)?" + RE_NS + rb"+(?:" + RE_S + rb"+[Aa](?:[Ss]|\xc5\xbf)" +
                     RE_S + rb"+" + RE_NS + rb"+)?" + RE_S + rb"*(?:#[^\n]+)?\Z")


class ParseError(Exception):
    pass


# ---- the pieces of Go's standard library the parser leans on -------------------------------------------------------
def decode_rune(b: bytes, i: int):
    """utf8.DecodeRune(b[i:]) -> (rune, width)"""
    c = b[i]
    if c < 0x80:
        return c, 1
    if 0xC2 <= c <= 0xDF:
        need, lo, hi, r = 1, 0x80, 0xBF, c & 0x1F
    elif 0xE0 <= c <= 0xEF:
        need, r = 2, c & 0x0F
        lo, hi = (0xA0 if c == 0xE0 else 0x80), (0x9F if c == 0xED else 0xBF)
    elif 0xF0 <= c <= 0xF4:
        need, r = 3, c & 0x07
        lo, hi = (0x90 if c == 0xF0 else 0x80), (0x8F if c == 0xF4 else 0xBF)
    else:
        return 0xFFFD, 1
    if i + need >= len(b):
        return 0xFFFD, 1
    c1 = b[i + 1]
    if not lo <= c1 <= hi:
        return 0xFFFD, 1
    r = (r << 6) | (c1 & 0x3F)
    for k in range(2, need + 1):
        ck = b[i + k]
        if not 0x80 <= ck <= 0xBF:
            return 0xFFFD, 1
        r = (r << 6) | (ck & 0x3F)
    return r, need + 1


_SPACE = {0x09, 0x0A, 0x0B, 0x0C, 0x0D, 0x20, 0x85, 0xA0, 0x1680, 0x2028, 0x2029, 0x202F, 0x205F, 0x3000} | set(range(0x2000, 0x200B))


def is_space(r: int) -> bool:  # unicode.IsSpace
    return r in _SPACE


def trim_left_space(b: bytes) -> bytes:  # bytes.TrimLeftFunc(b, unicode.IsSpace)
    i = 0
    while i < len(b):
        r, w = decode_rune(b, i)
        if not is_space(r):
            break
        i += w
    return b[i:]


def trim_space(b: bytes) -> bytes:  # strings.TrimSpace
    b = trim_left_space(b)
    end = len(b)
    while end > 0:
        # utf8.DecodeLastRune: step back over at most 3 continuation bytes to a start byte that decodes to exactly here
        start = end - 1
        lim = max(0, end - 4)
        r, w = 0xFFFD, 1
        k = start
        while k >= lim:
            if b[k] < 0x80 or b[k] >= 0xC0:
                rr, ww = decode_rune(b[:end], k)
                if k + ww == end:
                    r, w = rr, ww
                break
            k -= 1
        if not is_space(r):
            break
        end -= w
    return b[:end]


def to_lower(b: bytes) -> bytes:  # strings.ToLower
    if all(c < 0x80 for c in b):
        return b.lower()
    out = []
    i = 0
    while i < len(b):
        r, w = decode_rune(b, i)
        i += w
        if r < 0x80:
            out.append(chr(r).lower())
        elif 0xD800 <= r <= 0xDFFF:
            out.append("�")
        else:
            out.append(chr(r).lower()[0])  # Go maps rune to rune: U+0130 -> 'i' (Python would add a combining dot)
    return "".join(out).encode("utf-8")


# ---- parser.go -----------------------------------------------------------------------------------------------------
class Directive:
    def __init__(self):
        self.escape_token = ord("\\")
        self.line_continuation = re.compile(rb"\\[ \t]*\Z")
        self.processing_complete = False
        self.escape_seen = False

    def set_escape_token(self, s: bytes):
        if s not in (b"`", b"\\"):
            raise ParseError("invalid ESCAPE")
        self.escape_token = s[0]
        self.line_continuation = re.compile(re.escape(s) + rb"[ \t]*\Z")

    def possible_parser_directive(self, line: bytes):
        if self.processing_complete:
            return
        m = TOKEN_ESCAPE_COMMAND.match(to_lower(line))
        if m:
            if self.escape_seen:
                raise ParseError("only one escape parser directive can be used")
            self.escape_seen = True
            self.set_escape_token(m.group("escapechar"))
            return
        self.processing_complete = True


def trim_comments(src: bytes) -> bytes:
    return TOKEN_COMMENT.sub(b"", src)


def is_comment(line: bytes) -> bool:
    return TOKEN_COMMENT.match(trim_left_space(line)) is not None


def is_empty_continuation_line(line: bytes) -> bool:
    return len(trim_left_space(line)) == 0


def trim_continuation_character(line: bytes, d: Directive):
    if d.line_continuation.search(line):
        return d.line_continuation.sub(b"", line), False
    return line, True


def process_line(d: Directive, token: bytes, strip_left_whitespace: bool) -> bytes:
    if strip_left_whitespace:
        token = trim_left_space(token)
    d.possible_parser_directive(token)
    return trim_comments(token)


def scan_lines(data: bytes):
    """bufio.Scanner with ScanLines and the default 64 KiB buffer: yields lines, raises when one does not fit"""
    pos, n = 0, len(data)
    while pos < n:
        nl = data.find(b"\n", pos)
        end = n if nl < 0 else nl
        if end - pos >= MAX_SCAN_TOKEN:
            raise ParseError("dockerfile line greater than max allowed size")
        line = data[pos:end]
        if line.endswith(b"\r"):
            line = line[:-1]
        yield line
        pos = end + 1


def parse(data: bytes):
    """parser.Parse: list of (value, original) for the top-level nodes"""
    d = Directive()
    nodes = []
    lines = scan_lines(data)
    current_line = 0
    for raw in lines:
        if current_line == 0 and raw.startswith(b"\xef\xbb\xbf"):
            raw = raw[3:]
        bytes_read = process_line(d, raw, True)
        current_line += 1
        line, is_end_of_line = trim_continuation_character(bytes_read, d)
        if is_end_of_line and line == b"":
            continue
        while not is_end_of_line:
            try:
                raw = next(lines)
            except StopIteration:
                break
            bytes_read = process_line(d, raw, False)
            current_line += 1
            if is_comment(raw):
                continue
            if is_empty_continuation_line(bytes_read):
                continue
            continuation_line, is_end_of_line = trim_continuation_character(bytes_read, d)
            line += continuation_line
        nodes.append((new_node_from_line(line, d), line))
    if not nodes:
        raise ParseError("file with no instructions.")
    return nodes


# ---- split_command.go ----------------------------------------------------------------------------------------------
def split_command(line: bytes):
    cmdline = TOKEN_WHITESPACE.split(trim_space(line), 1)
    cmd = to_lower(cmdline[0])
    args = b""
    if len(cmdline) == 2:
        args = extract_builder_flags(cmdline[1])
    return cmd, trim_space(args)


def extract_builder_flags(line: bytes) -> bytes:
    """returns the remaining part of the line (the flags themselves never make Parse fail)"""
    IN_SPACES, IN_WORD, IN_QUOTE = 0, 1, 2
    phase = IN_SPACES
    word = ""
    quote = 0
    ch = 0
    pos = 0
    n = len(line)
    while pos <= n:
        if pos != n:
            ch = line[pos]  # rune(line[pos]): a BYTE widened to a rune
        if phase == IN_SPACES:
            if pos == n:
                break
            if is_space(ch):
                pos += 1
                continue
            if ch != ord("-") or pos + 1 == n or line[pos + 1] != ord("-"):
                return line[pos:]
            phase = IN_WORD
        if phase in (IN_WORD, IN_QUOTE) and pos == n:
            break
        if phase == IN_WORD:
            if is_space(ch):
                phase = IN_SPACES
                if word == "--":
                    return line[pos:]
                word = ""
                pos += 1
                continue
            if ch in (ord("'"), ord('"')):
                quote = ch
                phase = IN_QUOTE
                pos += 1
                continue
            if ch == ord("\\"):
                if pos + 1 == n:
                    pos += 1
                    continue
                pos += 1
                ch = line[pos]
            word += chr(ch)
            pos += 1
            continue
        if phase == IN_QUOTE:
            if ch == quote:
                phase = IN_WORD
                pos += 1
                continue
            if ch == ord("\\"):
                if pos + 1 == n:
                    phase = IN_WORD
                    pos += 1
                    continue
                pos += 1
                ch = line[pos]
            word += chr(ch)
        pos += 1
    return b""


# ---- line_parsers.go -----------------------------------------------------------------------------------------------
def parse_words(rest: bytes, d: Directive):
    IN_SPACES, IN_WORD, IN_QUOTE = 0, 1, 2
    words = []
    phase = IN_SPACES
    word = []
    quote = 0
    blank_ok = False
    ch, w = 0, 0
    pos, n = 0, len(rest)
    while pos <= n:
        if pos != n:
            ch, w = decode_rune(rest, pos)
        if phase == IN_SPACES:
            if pos == n:
                break
            if is_space(ch):
                pos += w
                continue
            phase = IN_WORD
        if phase in (IN_WORD, IN_QUOTE) and pos == n:
            if blank_ok or word:
                words.append(word)
            break
        if phase == IN_WORD:
            if is_space(ch):
                phase = IN_SPACES
                if blank_ok or word:
                    words.append(word)
                word = []
                blank_ok = False
                pos += w
                continue
            if ch in (ord("'"), ord('"')):
                quote = ch
                blank_ok = True
                phase = IN_QUOTE
            if ch == d.escape_token:
                if pos + w == n:
                    pos += w
                    continue
                word.append(ch)
                pos += w
                ch, w = decode_rune(rest, pos)
            word.append(ch)
            pos += w
            continue
        if phase == IN_QUOTE:
            if ch == quote:
                phase = IN_WORD
            if ch == d.escape_token and quote != ord("'"):
                if pos + w == n:
                    phase = IN_WORD
                    pos += w
                    continue
                pos += w
                word.append(ch)
                ch, w = decode_rune(rest, pos)
            word.append(ch)
        pos += w
    return words


def parse_name_val(rest: bytes, d: Directive):
    words = parse_words(rest, d)
    if not words:
        return
    if ord("=") not in words[0]:
        parts = TOKEN_WHITESPACE.split(rest, 1)
        if len(parts) < 2:
            raise ParseError("must have two arguments")
        return
    for word in words:
        if ord("=") not in word:
            raise ParseError("can't find = in word")


class NotStringArray(ParseError):
    """errDockerfileNotStringArray: 'when using JSON array syntax, arrays must be comprised of strings only'"""


class JSONSyntax(Exception):
    """any other error of json.Decoder.Decode: parseMaybeJSON* then takes the argument as a plain string"""


MAX_NESTING_DEPTH = 10000  # encoding/json scanner.go (Go >= 1.15; the reference builds with ^1.15)
_JSON_SPACE = b" \t\r\n"
PARSE_OBJECT_KEY, PARSE_OBJECT_VALUE, PARSE_ARRAY_VALUE = 0, 1, 2


def json_decode_array(rest: bytes):
    """json.NewDecoder(strings.NewReader(rest)).Decode(&[]interface{}) for rest starting with '[': the FIRST value of the
    stream (what follows it is not looked at), by encoding/json's scanner state machine, state by state.  Returns, per
    top-level element, whether it is a string.  Raises JSONSyntax for whatever makes Decode return an error."""
    parse_state = []
    is_string = []       # per top-level element
    numbers = []         # literal of every number, at any depth: each goes through strconv.ParseFloat
    step = "beginValue"
    i, n = 0, len(rest)
    num_start = -1

    def push(kind):
        parse_state.append(kind)
        if len(parse_state) > MAX_NESTING_DEPTH:
            raise JSONSyntax("exceeded max depth")

    while True:
        if i >= n:
            raise JSONSyntax("unexpected EOF")  # io.ErrUnexpectedEOF: the array never closed
        c = rest[i]
        redo = False  # the state hands the same byte on to stateEndValue
        if step in ("beginValue", "beginValueOrEmpty", "beginStringOrEmpty", "beginString"):
            if c in _JSON_SPACE:
                i += 1
                continue
            if step == "beginValueOrEmpty" and c == ord("]"):
                step, redo = "endValue", True
            elif step == "beginStringOrEmpty" and c == ord("}"):
                parse_state[-1] = PARSE_OBJECT_VALUE
                step, redo = "endValue", True
            elif step in ("beginStringOrEmpty", "beginString"):
                if c != ord('"'):
                    raise JSONSyntax("looking for beginning of object key string")
                step = "inString"
            else:  # stateBeginValue
                if len(parse_state) == 1 and parse_state[0] == PARSE_ARRAY_VALUE:
                    is_string.append(c == ord('"'))
                if c == ord("{"):
                    push(PARSE_OBJECT_KEY)
                    step = "beginStringOrEmpty"
                elif c == ord("["):
                    push(PARSE_ARRAY_VALUE)
                    step = "beginValueOrEmpty"
                elif c == ord('"'):
                    step = "inString"
                elif c == ord("-"):
                    step, num_start = "neg", i
                elif c == ord("0"):
                    step, num_start = "0", i
                elif ord("1") <= c <= ord("9"):
                    step, num_start = "1", i
                elif c in b"tfn":
                    word = {ord("t"): b"true", ord("f"): b"false", ord("n"): b"null"}[c]
                    if rest[i:i + len(word)] != word:
                        raise JSONSyntax("in literal")
                    i += len(word) - 1
                    step = "endValue"
                else:
                    raise JSONSyntax("looking for beginning of value")
        elif step == "inString":
            if c == ord('"'):
                step = "endValue"
            elif c == ord("\\"):
                step = "inStringEsc"
            elif c < 0x20:
                raise JSONSyntax("in string literal")
        elif step == "inStringEsc":
            if c in b'bfnrt\\/"':
                step = "inString"
            elif c == ord("u"):
                hexd = rest[i + 1:i + 5]
                if len(hexd) < 4 or any(h not in b"0123456789abcdefABCDEF" for h in hexd):
                    raise JSONSyntax("in \\u hexadecimal character escape")
                i += 4
                step = "inString"
            else:
                raise JSONSyntax("in string escape code")
        elif step == "neg":
            if c == ord("0"):
                step = "0"
            elif ord("1") <= c <= ord("9"):
                step = "1"
            else:
                raise JSONSyntax("in numeric literal")
        elif step == "1":
            if not ord("0") <= c <= ord("9"):
                step, redo = "0", True
        elif step == "0":
            if c == ord("."):
                step = "dot"
            elif c in b"eE":
                step = "E"
            else:
                numbers.append(rest[num_start:i])
                step, redo = "endValue", True
        elif step == "dot":
            if not ord("0") <= c <= ord("9"):
                raise JSONSyntax("after decimal point in numeric literal")
            step = "dot0"
        elif step == "dot0":
            if ord("0") <= c <= ord("9"):
                pass
            elif c in b"eE":
                step = "E"
            else:
                numbers.append(rest[num_start:i])
                step, redo = "endValue", True
        elif step == "E":
            if c in b"+-":
                step = "ESign"
            else:
                step, redo = "ESign", True
        elif step == "ESign":
            if not ord("0") <= c <= ord("9"):
                raise JSONSyntax("in exponent of numeric literal")
            step = "E0"
        elif step == "E0":
            if not ord("0") <= c <= ord("9"):
                numbers.append(rest[num_start:i])
                step, redo = "endValue", True
        elif step == "endValue":
            if c in _JSON_SPACE:
                pass
            else:
                ps = parse_state[-1]
                if ps == PARSE_OBJECT_KEY:
                    if c != ord(":"):
                        raise JSONSyntax("after object key")
                    parse_state[-1] = PARSE_OBJECT_VALUE
                    step = "beginValue"
                elif ps == PARSE_OBJECT_VALUE:
                    if c == ord(","):
                        parse_state[-1] = PARSE_OBJECT_KEY
                        step = "beginString"
                    elif c == ord("}"):
                        parse_state.pop()
                    else:
                        raise JSONSyntax("after object key:value pair")
                else:
                    if c == ord(","):
                        step = "beginValue"
                    elif c == ord("]"):
                        parse_state.pop()
                        if not parse_state:
                            break  # the top-level array closed: the value is complete
                    else:
                        raise JSONSyntax("after array element")
        if not redo:
            i += 1
    for lit in numbers:  # d.convertNumber: strconv.ParseFloat(s, 64) fails on overflow only
        if float(lit) in (float("inf"), float("-inf")):
            raise JSONSyntax("UnmarshalTypeError: number out of range")
    return is_string


def parse_json(rest: bytes):
    rest = trim_left_space(rest)
    if not rest.startswith(b"["):
        raise JSONSyntax("Error parsing as a JSON array")
    for s in json_decode_array(rest):
        if not s:
            raise NotStringArray()


def parse_maybe_json(rest: bytes):
    """parseMaybeJSON and parseMaybeJSONToList agree on what makes them fail"""
    if rest == b"":
        return
    try:
        parse_json(rest)
    except JSONSyntax:
        return  # taken as a plain string / a whitespace-delimited list


def parse_health_config(rest: bytes):
    sep = 0
    while sep < len(rest) and not is_space(rest[sep]):  # unicode.IsSpace(rune(rest[sep])): byte by byte
        sep += 1
    nxt = sep
    while nxt < len(rest) and is_space(rest[nxt]):
        nxt += 1
    if sep == 0:
        return
    parse_maybe_json(rest[nxt:])


JSON_COMMANDS = (b"add", b"cmd", b"copy", b"entrypoint", b"run", b"shell", b"volume")


def new_node_from_line(line: bytes, d: Directive) -> bytes:
    """returns the node's Value (the lower-cased command); raises on what makes newNodeFromLine fail"""
    cmd, args = split_command(line)
    if cmd in (b"env", b"label"):
        parse_name_val(args, d)
    elif cmd == b"onbuild":
        if args != b"":
            new_node_from_line(args, d)  # parseSubCommand
    elif cmd in JSON_COMMANDS:
        parse_maybe_json(args)
    elif cmd == b"healthcheck":
        parse_health_config(args)
    # the others cannot fail: parseIgnore (unknown commands), parseString (maintainer, stopsignal, user, workdir),
    # parseStringsWhitespaceDelimited (expose, from), parseNameOrNameVal (arg)
    return cmd


# ---- dockerfile2kube.go:117-144 ------------------------------------------------------------------------------------
def is_dockerfile(data: bytes) -> bool:
    try:
        nodes = parse(data)
    except ParseError:
        return False
    for value, original in nodes:
        if value == b"from":
            return FROM_RE.search(original) is not None
        if value == b"arg":
            continue
        return False
    return False
