// Device-side isDockerFile (reference: internal/source/dockerfile2kube.go:117-144) -- a streaming,
// allocation-free restatement of moby/buildkit v0.7.2 frontend/dockerfile/parser.Parse followed by
// the reference's "first non-ARG instruction is FROM and matches the regexp" rule.
//
// [external] buildkit's source is not in the reference tree (go.mod:22 pins v0.7.2) and the
// reference has no test for this function: parity for this kernel is unpinned.  What
// Parse can fail on, and therefore what this code checks for the whole file: the escape parser
// directive (invalid or repeated), ENV/LABEL name-value syntax (also under ONBUILD), "file with no
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

__device__ inline bool all_space(const File& f, uint64_t lb, uint64_t le) {
  while (lb < le) {
    uint32_t w, r = decode_rune(f.p + lb, le - lb, &w);
    if (!is_space_rune(r)) return false;
    lb += w;
  }
  return true;
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
    if (lb < le && f.p[lb] == '#') continue;  // isComment(raw continuation line)
    if (all_space(f, lb, le)) continue;       // isEmptyContinuationLine
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

// keywords: bit0 from, bit1 arg, bit2 env, bit3 label, bit4 onbuild (packed little-endian)
__device__ __forceinline__ uint32_t kw_step(uint32_t alive, uint32_t idx, uint32_t r) {
  // strings.ToLower: the only non-ASCII runes lower-casing into ASCII letters are U+0130 and U+212A
  if (r - 'A' < 26u) r += 32; else if (r == 0x130) r = 'i'; else if (r == 0x212A) r = 'k';
  constexpr uint64_t KW[5] = {0x6D6F7266ull /*from*/, 0x677261ull /*arg*/, 0x766E65ull /*env*/,
                              0x6C6562616Cull /*label*/, 0x646C6975626E6Full /*onbuild*/};
  if (idx >= 8) return 0;
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 5; k++)
    if ((alive >> k & 1) && (uint32_t)((KW[k] >> (8 * idx)) & 0xFF) == r) out |= 1u << k;
  return out;
}
__device__ __forceinline__ uint32_t kw_final(uint32_t alive, uint32_t len) {
  constexpr uint32_t LEN[5] = {4, 3, 3, 5, 7};
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 5; k++) if ((alive >> k & 1) && len == LEN[k]) out |= 1u << k;
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

// newNodeFromLine at c.  Returns false on a parse error; *kind: 0 other, 1 from, 2 arg.
__device__ inline bool node_from_line(const File& f, LCur c, int* kind) {
  *kind = 0;
  for (int depth = 0;; depth++) {
    // strings.TrimSpace (left)
    for (;;) { LCur t = c; uint32_t r; if (!lc_get_rune(f, t, &r)) { return true; } if (!is_space_rune(r)) break; c = t; }
    uint32_t alive = 0x1F, idx = 0;
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
    if (depth == 0) *kind = (kw & 1) ? 1 : (kw & 2) ? 2 : 0;
    if (!(kw & (4 | 8 | 16))) return true;  // only ENV, LABEL and ONBUILD can fail
    if (!has_rest) return true;             // no arguments: parseNameVal / parseSubCommand accept ""
    for (;;) { LCur t = c; int b = lc_get(f, t); if (b < 0 || !ascii_ws((uint32_t)b)) break; c = t; }
    skip_builder_flags(f, c);
    if (kw & (4 | 8)) return name_val_ok(f, c);
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
  // from IN: comment and blank lines are dropped and the continuation goes on
  bool cont = false;
  { uint64_t e = le; cont = trim_continuation(f, lb, &e); }
  bool raw_comment = lb < le && f.p[lb] == '#';
  bool blank_line = all_space(f, lb, le);
  uint32_t from_in = (raw_comment || blank_line) ? 1u : (cont ? 1u : 0u);
  // from OUT: the line is left-trimmed first; comments and empty lines are skipped
  uint64_t t = lb;
  while (t < le) { uint32_t w, r = decode_rune(f.p + t, le - t, &w); if (!is_space_rune(r)) break; t += w; }
  uint32_t from_out = (t == le || f.p[t] == '#') ? 0u : (cont ? 1u : 0u);
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
        // only ENV, LABEL and ONBUILD can fail: skip the line when its first word, ended by a blank
        // inside this physical line, cannot be one of them
        uint32_t alive = 4 | 8 | 16, idx = 0;
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
