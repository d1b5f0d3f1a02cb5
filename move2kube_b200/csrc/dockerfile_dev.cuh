// Device-side isDockerFile (reference: internal/source/dockerfile2kube.go:117-144) -- a streaming,
// allocation-free restatement of moby/buildkit v0.7.2 frontend/dockerfile/parser.Parse followed by
// the reference's "first non-ARG instruction is FROM and matches the regexp" rule.
//
// [external] buildkit's source is not in the reference tree (go.mod:22 pins v0.7.2) and the
// reference has no test for this function: parity for this kernel is unpinned.  What
// Parse can fail on, and therefore what this code checks for the whole file: the escape parser
// directive (invalid or repeated), ENV/LABEL name-value syntax and a JSON-form argument that is a well-formed array with a
// non-string element (both also under ONBUILD), "file with no
// instructions", and lines of 64 KiB or more (checked by the byte-scan kernel, passed in here).
//
// One thread parses one file.  A logical line (continuation lines joined, comment and blank
// continuation lines dropped) is never materialised: `LCur` walks it in place and is cheap to
// copy, which is how the regexp's alternatives are tried.
#pragma once
#include <stdint.h>
#include "gostr.h"

namespace m2k {
namespace df {

struct File {
  const uint8_t* p;
  uint64_t n;
  uint32_t esc;  // current escape token
};

__device__ __forceinline__ bool blank(uint32_t c) { return c == ' ' || c == '\t'; }
__device__ __forceinline__ bool ascii_ws(uint32_t c) { return c == '\t' || c == '\v' || c == '\f' || c == '\r' || c == ' '; }  // tokenWhitespace
__device__ __forceinline__ bool re_s(uint32_t c) { return c == '\t' || c == '\n' || c == '\f' || c == '\r' || c == ' '; }        // RE2 \s

// physical line starting at pos: [lb, le) without '\n' and one trailing '\r'; returns start of next line.
// File bases are 16-byte aligned and zero padded to 16 bytes (include/m2kdetect.h), so the search for
// the newline reads aligned 8-byte words (SWAR) once pos is aligned.
__device__ inline uint64_t phys_line(const File& f, uint64_t pos, uint64_t* le) {
  uint64_t e = pos;
  while (e < f.n && (e & 7) && f.p[e] != '\n') e++;
  if (e < f.n && !(e & 7) && f.p[e] != '\n') {
    const uint64_t ones = 0x0101010101010101ull, nl = 0x0A0A0A0A0A0A0A0Aull;
    for (;;) {
      uint64_t w = *reinterpret_cast<const uint64_t*>(f.p + e) ^ nl;
      uint64_t z = (w - ones) & ~w & 0x8080808080808080ull;
      if (z) { e += (uint64_t)(__ffsll((long long)z) - 1) >> 3; break; }
      e += 8;
      if (e >= f.n) break;
    }
    if (e > f.n) e = f.n;
  }
  uint64_t next = e + 1;
  if (e > pos && f.p[e - 1] == '\r') e--;
  *le = e;
  return next;
}

// trimContinuationCharacter on [lb, le): regexp `\<esc>[ \t]*$`
__device__ inline bool trim_continuation(const File& f, uint64_t lb, uint64_t* le) {
  uint64_t e = *le;
  while (e > lb && blank(f.p[e - 1])) e--;
  if (e > lb && f.p[e - 1] == f.esc) { *le = e - 1; return true; }
  return false;
}

// cursor over a logical line
struct LCur {
  uint64_t pos, seg_end, next_line;
  bool more;  // the current segment ended with a continuation marker
};

__device__ inline int lc_get(const File& f, LCur& c) {
  for (;;) {
    if (c.pos < c.seg_end) return f.p[c.pos++];
    if (!c.more) return -1;
    if (c.next_line >= f.n) { c.more = false; return -1; }
    uint64_t lb = c.next_line, le;
    c.next_line = phys_line(f, lb, &le);
    // isComment(trimWhitespace(raw continuation line)) -- an indented comment is dropped too -- and isEmptyContinuationLine
    uint64_t t = lb;
    while (t < le) { uint32_t w, r = decode_rune(f.p + t, le - t, &w); if (!is_space_rune(r)) break; t += w; }
    if (t == le || f.p[t] == '#') continue;
    c.more = trim_continuation(f, lb, &le);
    c.pos = lb;
    c.seg_end = le;
  }
}
__device__ inline int lc_peek(const File& f, const LCur& c) { LCur t = c; return lc_get(f, t); }

// next rune of the joined line (utf8.DecodeRuneInString semantics across segment joins)
__device__ inline bool lc_get_rune(const File& f, LCur& c, uint32_t* r) {
  int b0 = lc_get(f, c);
  if (b0 < 0) return false;
  if (b0 < 0x80) { *r = (uint32_t)b0; return true; }
  uint8_t buf[4];
  buf[0] = (uint8_t)b0;
  LCur t = c;
  uint32_t have = 1;
  while (have < 4) { int b = lc_get(f, t); if (b < 0) break; buf[have++] = (uint8_t)b; }
  uint32_t w;
  *r = decode_rune(buf, have, &w);
  for (uint32_t i = 1; i < w; i++) lc_get(f, c);
  return true;
}

__device__ inline bool lc_rest_all_space(const File& f, LCur c) {
  uint32_t r;
  while (lc_get_rune(f, c, &r)) if (!is_space_rune(r)) return false;
  return true;
}

// keywords: bit0 from, bit1 arg, bit2 env, bit3 label, bit4 onbuild -- the five the head of a file is about -- then
// the commands whose arguments may be a JSON array (parseMaybeJSON: cmd, entrypoint, run, shell; parseMaybeJSONToList:
// add, copy, volume; parseHealthConfig: healthcheck).  Two little-endian words of eight characters per keyword.
enum : uint32_t {
  KW_FROM = 1u << 0, KW_ARG = 1u << 1, KW_ENV = 1u << 2, KW_LABEL = 1u << 3, KW_ONBUILD = 1u << 4,
  KW_ADD = 1u << 5, KW_CMD = 1u << 6, KW_COPY = 1u << 7, KW_ENTRYPOINT = 1u << 8, KW_RUN = 1u << 9, KW_SHELL = 1u << 10,
  KW_VOLUME = 1u << 11, KW_HEALTHCHECK = 1u << 12,
  KW_HEAD = 0x1Fu,                                  // what first_word_rules_out and the head's quick reject follow
  KW_JSON = 0x1FE0u,                                // the eight JSON-form commands
  KW_CAN_FAIL = KW_ENV | KW_LABEL | KW_ONBUILD | KW_JSON  // newNodeFromLine can return an error for these only
};
constexpr int kNKeywords = 13;
__device__ __forceinline__ uint32_t kw_step(uint32_t alive, uint32_t idx, uint32_t r) {
  // strings.ToLower: the only non-ASCII runes lower-casing into ASCII letters are U+0130 and U+212A
  if (r - 'A' < 26u) r += 32; else if (r == 0x130) r = 'i'; else if (r == 0x212A) r = 'k';
  constexpr uint64_t LO[kNKeywords] = {0x6D6F7266ull /*from*/, 0x677261ull /*arg*/, 0x766E65ull /*env*/, 0x6C6562616Cull /*label*/,
                                       0x646C6975626E6Full /*onbuild*/, 0x646461ull /*add*/, 0x646D63ull /*cmd*/, 0x79706F63ull /*copy*/,
                                       0x696F707972746E65ull /*entrypoi*/, 0x6E7572ull /*run*/, 0x6C6C656873ull /*shell*/,
                                       0x656D756C6F76ull /*volume*/, 0x686368746C616568ull /*healthch*/};
  constexpr uint32_t HI[kNKeywords] = {0, 0, 0, 0, 0, 0, 0, 0, 0x746Eu /*nt*/, 0, 0, 0, 0x6B6365u /*eck*/};
  if (idx >= 12) return 0;
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < kNKeywords; k++) {
    if (!(alive >> k & 1)) continue;
    const uint32_t ch = idx < 8 ? (uint32_t)((LO[k] >> (8 * idx)) & 0xFF) : ((HI[k] >> (8 * (idx - 8))) & 0xFF);
    if (ch == r) out |= 1u << k;
  }
  return out;
}
__device__ __forceinline__ uint32_t kw_final(uint32_t alive, uint32_t len) {
  constexpr uint32_t LEN[kNKeywords] = {4, 3, 3, 5, 7, 3, 3, 4, 10, 3, 5, 6, 11};
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < kNKeywords; k++) if ((alive >> k & 1) && len == LEN[k]) out |= 1u << k;
  return out;
}

// The byte scan's pre-filter (kernels.cu first_word_rules_out_dockerfile): w = the first 16 bytes of a file, little
// endian, zero padded.  true = the first instruction cannot be FROM, whatever follows.
__device__ inline bool first_word_rules_out(const uint32_t (&w)[4]) {
  int start = ((w[0] & 0xFFFFFFu) == 0xBFBBEFu) ? 3 : 0;  // UTF-8 byte order mark
  uint32_t alive = 0x1F, idx = 0;
  bool in_word = false;
#pragma unroll
  for (int i = 0; i < 16; i++) {
    if (i < start) continue;
    uint32_t c = (w[i >> 2] >> (8 * (i & 3))) & 0xFF;
    bool ws = c == ' ' || (c >= 0x09 && c <= 0x0D);
    if (!in_word) {
      if (ws) continue;
      if (c == 0 || c == '#' || c == '\\' || c >= 0x80) return false;
      in_word = true;
    } else {
      if (ws) return kw_final(alive, idx) == 0;
      if (c == 0 || c == '\\' || c >= 0x80) return false;
    }
    alive = kw_step(alive, idx, c);
    idx++;
    if (alive == 0) return true;
  }
  return false;
}

// extractBuilderFlags: advances c to the start of the remaining part.  Byte-wise, with
// unicode.IsSpace applied to byte values (0x85 and 0xA0 count as space).
__device__ inline void skip_builder_flags(const File& f, LCur& c) {
  enum { inSpaces, inWord, inQuote } phase = inSpaces;
  uint32_t wlen = 0, w0 = 0, w1 = 0, quote = 0;
  for (;;) {
    LCur here = c;
    int ch = lc_get(f, c);
    if (phase == inSpaces) {
      if (ch < 0) return;
      if (is_space_rune((uint32_t)ch)) continue;
      if (ch != '-' || lc_peek(f, c) != '-') { c = here; return; }
      phase = inWord; wlen = 0;
    }
    if (ch < 0) return;
    bool dashdash = (wlen == 2 && w0 == '-' && w1 == '-');
    if (phase == inWord) {
      if (is_space_rune((uint32_t)ch)) {
        phase = inSpaces;
        if (dashdash) { c = here; return; }
        wlen = 0;
        continue;
      }
      if (ch == '\'' || ch == '"') { quote = (uint32_t)ch; phase = inQuote; continue; }
      if (ch == '\\') {
        int nx = lc_get(f, c);
        if (nx < 0) return;  // "\" at the very end is skipped
        ch = nx;
      }
      if (wlen == 0) w0 = (uint32_t)ch; else if (wlen == 1) w1 = (uint32_t)ch;
      wlen++;
      continue;
    }
    // inQuote
    if ((uint32_t)ch == quote) { phase = inWord; continue; }
    if (ch == '\\') {
      int nx = lc_get(f, c);
      if (nx < 0) return;
      ch = nx;
    }
    if (wlen == 0) w0 = (uint32_t)ch; else if (wlen == 1) w1 = (uint32_t)ch;
    wlen++;
  }
}

// parseNameVal (ENV / LABEL) over the rest of the line at c.  true = no error.
__device__ inline bool name_val_ok(const File& f, LCur c) {
  // strings.TrimSpace(left)
  for (;;) { LCur t = c; uint32_t r; if (!lc_get_rune(f, t, &r)) return true; if (!is_space_rune(r)) break; c = t; }
  enum { inSpaces, inWord, inQuote } phase = inSpaces;
  uint32_t words = 0, wlen = 0, quote = 0;
  bool weq = false, blankOK = false, first_eq = false, all_eq = true;
  bool interior_ws = false, pending_ws = false, seen_nonspace = false;
  for (;;) {
    uint32_t ch = 0;
    bool have = lc_get_rune(f, c, &ch);
    if (have) {  // bookkeeping for tokenWhitespace.Split(TrimSpace(rest), 2)
      if (is_space_rune(ch)) { if (ascii_ws(ch) && seen_nonspace) pending_ws = true; }
      else { if (pending_ws) interior_ws = true; seen_nonspace = true; }
    }
    if (phase == inSpaces) {
      if (!have) break;
      if (is_space_rune(ch)) continue;
      phase = inWord;
    }
    if (!have) {
      if (blankOK || wlen > 0) { if (words == 0) first_eq = weq; if (!weq) all_eq = false; words++; }
      break;
    }
    if (phase == inWord) {
      if (is_space_rune(ch)) {
        phase = inSpaces;
        if (blankOK || wlen > 0) { if (words == 0) first_eq = weq; if (!weq) all_eq = false; words++; }
        wlen = 0; weq = false; blankOK = false;
        continue;
      }
      if (ch == '\'' || ch == '"') { quote = ch; blankOK = true; phase = inQuote; }
      if (ch == f.esc) {
        uint32_t nx;
        LCur t = c;
        if (!lc_get_rune(f, t, &nx)) continue;  // escape token at the end of the line: skipped
        // Go's check is against the TrimSpace'd rest: an escape followed only by spaces is also "at the end"
        if (is_space_rune(nx) && lc_rest_all_space(f, t)) { c = t; while (lc_get(f, c) >= 0) {} continue; }
        wlen++; c = t; ch = nx;
        if (is_space_rune(nx)) { if (ascii_ws(nx) && seen_nonspace) pending_ws = true; }
        else { if (pending_ws) interior_ws = true; seen_nonspace = true; }
      }
      wlen++;
      if (ch == '=') weq = true;
      continue;
    }
    // inQuote
    if (ch == quote) phase = inWord;
    if (ch == f.esc && quote != '\'') {
      uint32_t nx;
      LCur t = c;
      if (!lc_get_rune(f, t, &nx)) { phase = inWord; continue; }
      if (is_space_rune(nx) && lc_rest_all_space(f, t)) { phase = inWord; c = t; while (lc_get(f, c) >= 0) {} continue; }
      wlen++; c = t; ch = nx;
      if (is_space_rune(nx)) { if (ascii_ws(nx) && seen_nonspace) pending_ws = true; }
      else { if (pending_ws) interior_ws = true; seen_nonspace = true; }
    }
    wlen++;
    if (ch == '=') weq = true;
  }
  if (words == 0) return true;
  if (!first_eq) return interior_ws;  // old form "KEY name value" needs a blank-separated value
  return all_eq;
}

// \s*(#.+)?$
__device__ inline bool tail_ok(const File& f, LCur c) {
  int b;
  do { b = lc_get(f, c); } while (b >= 0 && re_s((uint32_t)b));
  if (b < 0) return true;
  return b == '#' && lc_peek(f, c) >= 0;
}

// [^\s]+ starting at c (first byte known non-space), then the rest of the pattern `rest_kind`:
// returns true if some way of ending the run leads to a match.  On return *after is the cursor at
// the end of the maximal run.
__device__ inline bool run_then_tail(const File& f, LCur c, LCur* after) {
  uint32_t idx = 0;
  for (;;) {
    LCur here = c;
    int b = lc_get(f, c);
    if (b < 0 || re_s((uint32_t)b)) { *after = here; return tail_ok(f, here); }
    if (b == '#' && idx >= 1 && lc_peek(f, c) >= 0) { *after = here; return true; }  // cut the run, `#.+` takes the rest
    idx++;
  }
}

// regexp (?i)FROM\s+(--platform=[^\s]+)?[^\s]+(\s+AS\s+[^\s]+)?\s*(#.+)?$  -- c is just past "FROM".
// The optional platform group adds nothing: with or without it the next token is one run of
// non-space characters (the pattern has no \s+ after the group).
__device__ inline bool from_match_here(const File& f, LCur c) {
  int b = lc_get(f, c);
  if (b < 0 || !re_s((uint32_t)b)) return false;
  for (;;) { LCur t = c; b = lc_get(f, t); if (b < 0) return false; if (!re_s((uint32_t)b)) break; c = t; }
  LCur after;
  if (run_then_tail(f, c, &after)) return true;
  // (\s+AS\s+[^\s]+)
  c = after;
  uint32_t nws = 0;
  for (;;) { LCur t = c; b = lc_get(f, t); if (b < 0) return false; if (!re_s((uint32_t)b)) break; c = t; nws++; }
  if (nws == 0) return false;
  b = lc_get(f, c);
  if (b != 'a' && b != 'A') return false;
  b = lc_get(f, c);
  if (b == 0xC5) { if (lc_get(f, c) != 0xBF) return false; }  // U+017F folds to 's' under (?i)
  else if (b != 's' && b != 'S') return false;
  nws = 0;
  for (;;) { LCur t = c; b = lc_get(f, t); if (b < 0) return false; if (!re_s((uint32_t)b)) break; c = t; nws++; }
  if (nws == 0) return false;
  return run_then_tail(f, c, &after);
}

__device__ inline bool from_regex(const File& f, LCur c) {
  uint32_t win = 0;
  for (;;) {
    int b = lc_get(f, c);
    if (b < 0) return false;
    uint32_t lc = (uint32_t)b;
    if (lc - 'A' < 26u) lc += 32;
    win = (win << 8) | lc;
    if (win == 0x66726F6Du /* "from" */ && from_match_here(f, c)) return true;
  }
}

// ---- arguments in JSON form ---------------------------------------------------------------------------------------
// parseMaybeJSON / parseMaybeJSONToList (line_parsers.go) hand an argument that starts with '[' to
// json.NewDecoder(..).Decode(&[]interface{}) and fail -- errDockerfileNotStringArray, which fails Parse -- only when
// that succeeds AND some element of the array is not a string; any Decode error makes them take the argument as a
// plain string instead.  So the exact question is: is the FIRST JSON value of the stream a well-formed array (by
// encoding/json's scanner: nesting at most 10000, Go >= 1.15), with every number at any depth inside float64 range
// (d.convertNumber: strconv.ParseFloat fails on overflow only), and with a top-level element that is not a string?

// |value| >= 2^1024 - 2^970 rounds to infinity.  The literal is followed digit by digit: `nd` significant digits so
// far, `cmp` how their first 309 compare with the limit's, `point` where the decimal point sits relative to the first
// significant digit.
struct F64Range {
  uint32_t nd; int32_t cmp; int64_t point; bool seen_point;
  __device__ void start() { nd = 0; cmp = 0; point = 0; seen_point = false; }
  __device__ void digit(uint32_t d) {
    const char* const LIMIT =
        "1797693134862315807937289714053034150799341327100378269361737789804449682927647509466490179775872070963302864166928879109465555478519404026306574886715058206819089020007083836762738548458177115317644757302700698555713669596228429148198608349364752927190741684443655107043427115596995080930428801779041744977"
        "92";
    if (nd == 0 && d == 0) { if (seen_point) point--; return; }  // a leading zero: only moves the point
    if (nd < 309 && cmp == 0) { const uint32_t l = (uint32_t)(LIMIT[nd] - '0'); cmp = d < l ? -1 : (d > l ? 1 : 0); }
    nd++;
    if (!seen_point) point++;
  }
  __device__ void decimal_point() { seen_point = true; }
  // exp: the exponent's value (clamped by the caller), already signed
  __device__ bool overflows(int64_t exp) const {
    if (nd == 0) return false;  // zero
    const int64_t p = point + exp;  // value = 0.d1d2d3... x 10^p
    if (p != 309) return p > 309;
    return nd >= 309 ? cmp >= 0 : cmp > 0;  // fewer digits than the limit and equal so far: the limit's next digit is not 0
  }
};

#ifndef M2K_JSON_KEEP
#define M2K_JSON_KEEP 64  // innermost levels whose kind (array / object) is kept in a register; tests build with 2
#endif

// c: cursor at the '[' .  true = errDockerfileNotStringArray.
__device__ inline bool json_array_not_all_strings(const File& f, LCur c) {
  enum { BEGIN_VALUE, BEGIN_VALUE_OR_EMPTY, BEGIN_STRING_OR_EMPTY, BEGIN_STRING, IN_STRING, IN_ESC, NEG, N1, N0, DOT, DOT0, EXP, ESIGN, E0, END_VALUE };
  const LCur origin = c;
  uint64_t kinds = 0;        // bit i: the container i levels above the innermost is an object
  uint32_t depth = 0, known = 0, consumed = 0;
  bool top_key = false;      // the innermost container is an object waiting for ':' after a key
  bool non_string = false, out_of_range = false, exp_neg = false;
  int step = BEGIN_VALUE;
  int64_t exp = 0;
  F64Range num;
  num.start();
  int b = lc_get(f, c);
  consumed++;
  for (;;) {
    if (b < 0) return false;  // the array never closed
    const uint32_t ch = (uint32_t)b;
    const bool sp = ch == ' ' || ch == '\t' || ch == '\r' || ch == '\n';
    bool redo = false;  // the state hands this byte on to another one
    switch (step) {
      case BEGIN_VALUE: case BEGIN_VALUE_OR_EMPTY: case BEGIN_STRING_OR_EMPTY: case BEGIN_STRING:
        if (sp) break;
        if (step == BEGIN_VALUE_OR_EMPTY && ch == ']') { step = END_VALUE; redo = true; break; }
        if (step == BEGIN_STRING_OR_EMPTY && ch == '}') { top_key = false; step = END_VALUE; redo = true; break; }
        if (step == BEGIN_STRING_OR_EMPTY || step == BEGIN_STRING) { if (ch != '"') return false; step = IN_STRING; break; }
        if (depth == 1 && ch != '"') non_string = true;
        if (ch == '{' || ch == '[') {
          if (++depth > 10000) return false;  // "exceeded max depth"
          kinds = (kinds << 1) | (ch == '{' ? 1u : 0u);
          if (known < M2K_JSON_KEEP) known++;
          top_key = ch == '{';
          step = ch == '{' ? BEGIN_STRING_OR_EMPTY : BEGIN_VALUE_OR_EMPTY;
        } else if (ch == '"') step = IN_STRING;
        else if (ch == '-') { step = NEG; num.start(); }
        else if (ch == '0') { step = N0; num.start(); num.digit(0); }
        else if (ch - '1' < 9u) { step = N1; num.start(); num.digit(ch - '0'); }
        else if (ch == 't' || ch == 'f' || ch == 'n') {
          const char* w = ch == 't' ? "true" : ch == 'f' ? "false" : "null";
          for (int k = 1; w[k]; k++) { b = lc_get(f, c); consumed++; if (b != w[k]) return false; }
          step = END_VALUE;
        } else return false;
        break;
      case IN_STRING:
        if (ch == '"') step = END_VALUE; else if (ch == '\\') step = IN_ESC; else if (ch < 0x20) return false;
        break;
      case IN_ESC:
        if (ch == 'b' || ch == 'f' || ch == 'n' || ch == 'r' || ch == 't' || ch == '\\' || ch == '/' || ch == '"') step = IN_STRING;
        else if (ch == 'u') {
          for (int k = 0; k < 4; k++) {
            b = lc_get(f, c); consumed++;
            if (b < 0) return false;
            const uint32_t h = (uint32_t)b;
            if (!(h - '0' < 10u || (h | 0x20) - 'a' < 6u)) return false;
          }
          step = IN_STRING;
        } else return false;
        break;
      case NEG: if (ch == '0') { step = N0; num.digit(0); } else if (ch - '1' < 9u) { step = N1; num.digit(ch - '0'); } else return false; break;
      case N1: if (ch - '0' < 10u) num.digit(ch - '0'); else { step = N0; redo = true; } break;
      case N0:
        if (ch == '.') { step = DOT; num.decimal_point(); }
        else if (ch == 'e' || ch == 'E') { step = EXP; exp = 0; exp_neg = false; }
        else { if (num.overflows(0)) out_of_range = true; step = END_VALUE; redo = true; }
        break;
      case DOT: if (ch - '0' >= 10u) return false; num.digit(ch - '0'); step = DOT0; break;
      case DOT0:
        if (ch - '0' < 10u) num.digit(ch - '0');
        else if (ch == 'e' || ch == 'E') { step = EXP; exp = 0; exp_neg = false; }
        else { if (num.overflows(0)) out_of_range = true; step = END_VALUE; redo = true; }
        break;
      case EXP: if (ch == '+' || ch == '-') { exp_neg = ch == '-'; step = ESIGN; } else { step = ESIGN; redo = true; } break;
      case ESIGN: if (ch - '0' >= 10u) return false; exp = ch - '0'; step = E0; break;
      case E0:
        if (ch - '0' < 10u) { if (exp < 100000000) exp = exp * 10 + (ch - '0'); }
        else { if (num.overflows(exp_neg ? -exp : exp)) out_of_range = true; step = END_VALUE; redo = true; }
        break;
      case END_VALUE: {
        if (sp) break;
        const bool in_object = (kinds & 1u) != 0;
        if (in_object && top_key) { if (ch != ':') return false; top_key = false; step = BEGIN_VALUE; break; }
        if (ch == ',') { if (in_object) { top_key = true; step = BEGIN_STRING; } else step = BEGIN_VALUE; break; }
        if (ch != (in_object ? '}' : ']')) return false;
        if (--depth == 0) return non_string && !out_of_range;  // the top-level array closed: the value is complete
        kinds >>= 1; known--;
        top_key = false;  // an enclosing object is inside one of its values
        if (known == 0) {
          // the kinds of the enclosing levels fell out of the register: one more pass over what has been read,
          // noting the last container opened at each of the innermost levels <= depth (those are the open ones)
          LCur r = origin;
          uint32_t d = 0;
          bool in_str = false;
          uint64_t k2 = 0;
          for (uint32_t i = 0; i < consumed; i++) {
            const int x = lc_get(f, r);
            if (in_str) { if (x == '\\') { lc_get(f, r); i++; } else if (x == '"') in_str = false; continue; }
            if (x == '"') in_str = true;
            else if (x == '[' || x == '{') {
              d++;
              if (d <= depth && depth - d < M2K_JSON_KEEP) { const uint64_t bit = 1ull << (depth - d); k2 = x == '{' ? (k2 | bit) : (k2 & ~bit); }
            } else if (x == ']' || x == '}') d--;
          }
          kinds = k2;
          known = depth < M2K_JSON_KEEP ? depth : M2K_JSON_KEEP;
        }
        break;
      }
    }
    if (!redo) { b = lc_get(f, c); consumed++; }
  }
}

// parseMaybeJSON* on the rest of the line at c (already past the flags): false = they fail
__device__ inline bool maybe_json_ok(const File& f, LCur c) {
  for (;;) { LCur t = c; uint32_t r; if (!lc_get_rune(f, t, &r)) return true; if (!is_space_rune(r)) break; c = t; }  // TrimSpace / parseJSON's TrimLeftFunc
  if (lc_peek(f, c) != '[') return true;
  return !json_array_not_all_strings(f, c);
}

// parseHealthConfig: the first word byte by byte (unicode.IsSpace of the byte's value), then parseMaybeJSON
__device__ inline bool health_config_ok(const File& f, LCur c) {
  for (;;) { LCur t = c; uint32_t r; if (!lc_get_rune(f, t, &r)) return true; if (!is_space_rune(r)) break; c = t; }  // TrimSpace(args)
  uint32_t sep = 0;
  for (;; sep++) { LCur t = c; const int b = lc_get(f, t); if (b < 0 || is_space_rune((uint32_t)b)) break; c = t; }
  if (sep == 0) return true;
  for (;;) { LCur t = c; const int b = lc_get(f, t); if (b < 0 || !is_space_rune((uint32_t)b)) break; c = t; }
  return maybe_json_ok(f, c);
}

// newNodeFromLine at c.  Returns false on a parse error; *kind: 0 other, 1 from, 2 arg.
__device__ inline bool node_from_line(const File& f, LCur c, int* kind) {
  *kind = 0;
  for (int depth = 0;; depth++) {
    // strings.TrimSpace (left)
    for (;;) { LCur t = c; uint32_t r; if (!lc_get_rune(f, t, &r)) { return true; } if (!is_space_rune(r)) break; c = t; }
    uint32_t alive = (1u << kNKeywords) - 1, idx = 0;
    bool has_rest = false;
    for (;;) {
      LCur t = c;
      uint32_t r;
      if (!lc_get_rune(f, t, &r)) break;
      if (ascii_ws(r)) { has_rest = true; break; }
      if (is_space_rune(r) && lc_rest_all_space(f, t)) break;  // trailing space removed by TrimSpace
      alive = kw_step(alive, idx, r);
      idx++;
      c = t;
    }
    uint32_t kw = kw_final(alive, idx);
    if (depth == 0) *kind = (kw & KW_FROM) ? 1 : (kw & KW_ARG) ? 2 : 0;
    if (!(kw & KW_CAN_FAIL)) return true;   // the other parsers cannot fail (parseIgnore, parseString, parseStringsWhitespaceDelimited, parseNameOrNameVal)
    if (!has_rest) return true;             // no arguments: every one of them accepts ""
    for (;;) { LCur t = c; int b = lc_get(f, t); if (b < 0 || !ascii_ws((uint32_t)b)) break; c = t; }
    skip_builder_flags(f, c);
    if (kw & (KW_ENV | KW_LABEL)) return name_val_ok(f, c);
    if (kw & KW_HEALTHCHECK) return health_config_ok(f, c);
    if (kw & KW_JSON) return maybe_json_ok(f, c);
    // onbuild: parseSubCommand -> newNodeFromLine(rest)
  }
}

// possibleParserDirective on [lb, le).  Returns false on error.
__device__ inline bool directive_ok(File& f, uint64_t lb, uint64_t le, bool* processing_complete, bool* escape_seen) {
  if (*processing_complete) return true;
  const uint8_t* s = f.p;
  uint64_t i = lb;
  bool m = false;
  uint32_t ch = 0;
  if (i < le && s[i] == '#') {
    i++;
    while (i < le && blank(s[i])) i++;
    const char* kw = "escape";
    uint32_t k = 0;
    while (k < 6 && i + k < le && (uint32_t)(s[i + k] | 0x20) == (uint32_t)kw[k] && ((s[i + k] | 0x20) - 'a' < 26u)) k++;
    if (k == 6) {
      i += 6;
      while (i < le && blank(s[i])) i++;
      if (i < le && s[i] == '=') {
        i++;
        uint64_t k0 = i;
        while (i < le && blank(s[i])) i++;
        if (i < le) { m = true; ch = s[i]; }
        else if (i > k0) { m = true; ch = ' '; }  // the greedy [ \t]* hands a blank back to `.`
      }
    }
  }
  if (m) {
    if (*escape_seen) return false;
    *escape_seen = true;
    if (ch != '`' && ch != '\\') return false;
    f.esc = ch;
    return true;
  }
  *processing_complete = true;
  return true;
}

// Head of the file: parser directives, comments, ARG lines, up to the first other instruction.
// Returns true when that instruction is FROM and matches the reference's regexp; then *body is the
// offset of the first physical line after it and *esc the escape token in force.  Everything after
// `body` can no longer change the answer except by a parse error, which the striped validation
// below looks for in parallel.  `long_line`: the byte scan found a line of >= 65536 bytes (Parse
// fails; it also keeps this thread from walking such a line).
// `touched_end` (optional): set when the verdict rests on the last line present in [0, n) -- for a
// caller that holds only a prefix of the file, the verdict is then not final.
__device__ inline bool head(const uint8_t* p, uint64_t n, bool long_line, uint64_t* body, uint32_t* esc, bool* touched_end = nullptr) {
  if (long_line || n == 0) return false;
  bool at_end = false;
  struct Note { bool* dst; bool* src; __device__ ~Note() { if (dst) *dst = *src; } } note{touched_end, &at_end};
  File f{p, n, '\\'};
  bool processing_complete = false, escape_seen = false;
  uint64_t pos = 0;
  bool first_line = true;
  while (pos < n) {
    uint64_t lb = pos, le;
    // quick reject before looking for the end of the line: once the first word can no longer become a
    // keyword, "first instruction is not FROM" is settled
    {
      uint64_t q = lb;
      if (first_line && n - q >= 3 && p[q] == 0xEF && p[q + 1] == 0xBB && p[q + 2] == 0xBF) q += 3;
      while (q < n && p[q] != '\n') { uint32_t w, r = decode_rune(p + q, n - q, &w); if (!is_space_rune(r)) break; q += w; }
      if (q < n && p[q] != '\n' && p[q] != '#') {
        uint32_t alive = 0x1F, idx = 0;
        uint64_t t = q;
        while (t < n && alive) {
          uint32_t c0 = p[t];
          if (c0 == '\n' || ascii_ws(c0)) break;
          if (c0 == f.esc) { alive = 0x1F; break; }  // may be a line continuation inside the word: exact path
          if (c0 >= 0x80 && n - t < 4) { alive = 0x1F; break; }  // a rune that [0, n) may hold only in part (n can be a cut): exact path
          uint32_t w, r = decode_rune(p + t, n - t, &w);
          if (is_space_rune(r)) break;  // unicode space: leave it to the exact path
          alive = kw_step(alive, idx, r);
          idx++; t += w;
        }
        if (alive == 0) return false;  // the command word cannot become from/arg/env/label/onbuild any more
      }
    }
    pos = phys_line(f, lb, &le);
    if (pos >= n) at_end = true;
    if (first_line) { if (le - lb >= 3 && p[lb] == 0xEF && p[lb + 1] == 0xBB && p[lb + 2] == 0xBF) lb += 3; first_line = false; }
    while (lb < le) { uint32_t w, r = decode_rune(p + lb, le - lb, &w); if (!is_space_rune(r)) break; lb += w; }  // trimWhitespace
    if (!directive_ok(f, lb, le, &processing_complete, &escape_seen)) return false;
    if (lb < le && p[lb] == '#') le = lb;  // trimComments
    bool more = trim_continuation(f, lb, &le);
    if (!more && le == lb) continue;
    LCur start{lb, le, pos, more};
    LCur endc = start;
    while (lc_get(f, endc) >= 0) {}
    if (endc.next_line >= n) at_end = true;
    int kind;
    if (!node_from_line(f, start, &kind)) return false;
    pos = endc.next_line;
    if (kind == 2) continue;  // ARG: keep looking
    if (kind != 1 || !from_regex(f, start)) return false;
    *body = pos < n ? pos : n;
    *esc = f.esc;
    return true;
  }
  at_end = true;
  return false;  // file with no instructions (or only ARG)
}

// ---- striped validation of the body ---------------------------------------------------------------
// Whether a physical line starts a logical line depends on one bit of state carried from the line
// before it: OUT (0) or inside a continuation (1).  Each line maps that state to the next, so the
// state at any offset is a prefix composition of 2-bit maps: bit0 = next state from OUT, bit1 = next
// state from IN.  Stripes of the body compute their composed map, a segmented scan with a look-back across
// tiles turns the maps into the state at each stripe start (k_df_states), and every stripe then checks the logical
// lines that start inside it (k_df_validate).
#ifndef M2K_DF_STRIPE
#define M2K_DF_STRIPE 128
#endif
constexpr uint32_t kStripe = M2K_DF_STRIPE;   // == kDfStripe (m2k_internal.h)

__device__ inline uint32_t line_map(const File& f, uint64_t lb, uint64_t le) {
  bool cont = false;
  { uint64_t e = le; cont = trim_continuation(f, lb, &e); }
  uint64_t t = lb;  // the line is left-trimmed before it is looked at, inside a continuation as well as outside
  while (t < le) { uint32_t w, r = decode_rune(f.p + t, le - t, &w); if (!is_space_rune(r)) break; t += w; }
  const bool dropped = t == le || f.p[t] == '#';  // blank or comment
  // from IN: dropped lines leave the continuation open; from OUT: they are skipped
  const uint32_t from_in = dropped ? 1u : (cont ? 1u : 0u);
  const uint32_t from_out = dropped ? 0u : (cont ? 1u : 0u);
  return from_out | (from_in << 1);
}
__device__ __forceinline__ uint32_t map_apply(uint32_t m, uint32_t state) { return (m >> state) & 1u; }
__device__ __forceinline__ uint32_t map_compose(uint32_t first, uint32_t then) {  // then(first(x))
  return map_apply(then, map_apply(first, 0)) | (map_apply(then, map_apply(first, 1)) << 1);
}

// first physical-line start in [a, b) (a line starts at `body` or right after a '\n'); >= b when no line
// starts there -- the search never leaves the stripe, however long the line that runs through it
__device__ inline uint64_t first_line_start(const File& f, uint64_t body, uint64_t a, uint64_t b) {
  if (a <= body) return body;
  uint64_t q = a - 1;  // a '\n' at a - 1 makes `a` itself a line start
  const uint64_t end = b - 1 < f.n ? b - 1 : f.n;  // a '\n' at b - 1 starts a line at b: the next stripe's
  while (q < end && (q & 7) && f.p[q] != '\n') q++;
  if (q < end && !(q & 7) && f.p[q] != '\n') {
    const uint64_t ones = 0x0101010101010101ull, nl = 0x0A0A0A0A0A0A0A0Aull;
    for (;;) {  // aligned 8-byte words; the file is zero padded to 16 bytes, so a word never leaves it
      uint64_t w = *reinterpret_cast<const uint64_t*>(f.p + q) ^ nl;
      uint64_t z = (w - ones) & ~w & 0x8080808080808080ull;
      if (z) { q += (uint64_t)(__ffsll((long long)z) - 1) >> 3; break; }
      q += 8;
      if (q >= end) break;
    }
  }
  if (q >= end) return b;
  return q + 1;
}

// composed map of the lines that START in [a, b)
__device__ inline uint32_t stripe_map(const File& f, uint64_t body, uint64_t a, uint64_t b) {
  uint32_t m = 0x2;  // identity: OUT->OUT, IN->IN
  uint64_t q = first_line_start(f, body, a, b);
  while (q < b && q < f.n) {
    uint64_t le, next = phys_line(f, q, &le);
    m = map_compose(m, line_map(f, q, le));
    q = next;
  }
  return m;
}

// checks every logical line that starts in [a, b), given the state at the first line start; false = parse error
__device__ inline bool stripe_validate(const File& f, uint64_t body, uint64_t a, uint64_t b, uint32_t state) {
  uint64_t q = first_line_start(f, body, a, b);
  while (q < b && q < f.n) {
    uint64_t le, next = phys_line(f, q, &le);
    if (state == 0) {
      uint64_t lb = q;
      while (lb < le) { uint32_t w, r = decode_rune(f.p + lb, le - lb, &w); if (!is_space_rune(r)) break; lb += w; }
      bool check = lb < le && f.p[lb] != '#';
      if (check) {
        // only ENV, LABEL, ONBUILD and the JSON-form commands can fail: skip the line when its first word, ended by
        // a blank inside this physical line, cannot be one of them
        uint32_t alive = KW_CAN_FAIL, idx = 0;
        for (uint64_t t = lb; t < le && alive; idx++) {
          uint32_t c0 = f.p[t];
          if (ascii_ws(c0)) break;
          if (c0 == f.esc || c0 >= 0x80) { alive = 0x1F; break; }  // continuation or non-ASCII: exact path
          alive = kw_step(alive, idx, c0);
          t++;
        }
        check = alive != 0;
      }
      if (check) {
        uint64_t e2 = le;
        bool more = trim_continuation(f, lb, &e2);
        LCur start{lb, e2, next, more};
        int kind;
        if (!node_from_line(f, start, &kind)) return false;
      }
    }
    state = map_apply(line_map(f, q, le), state);
    q = next;
  }
  return true;
}

}  // namespace df
}  // namespace m2k
