// Go string semantics the detection path depends on, usable from host code and kernels.
// [external] Go 1.15 stdlib (unicode/utf8, unicode, strings): call sites in the reference are
// internal/common/utils.go:310 (strings.EqualFold) and internal/source/any2kube.go:165-176.
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef __CUDACC__
#define M2K_HD __host__ __device__ __forceinline__
#else
#define M2K_HD inline
#endif

namespace m2k {

// utf8.DecodeRune over s[0..n): invalid or truncated sequences decode as (U+FFFD, width 1).
M2K_HD uint32_t decode_rune(const uint8_t* s, size_t n, uint32_t* width) {
  uint32_t c = s[0];
  *width = 1;
  if (c < 0x80) return c;
  uint32_t need, r, lo = 0x80, hi = 0xBF;
  if (c >= 0xC2 && c <= 0xDF) { need = 1; r = c & 0x1F; }
  else if ((c & 0xF0) == 0xE0) { need = 2; r = c & 0x0F; if (c == 0xE0) lo = 0xA0; if (c == 0xED) hi = 0x9F; }
  else if (c >= 0xF0 && c <= 0xF4) { need = 3; r = c & 0x07; if (c == 0xF0) lo = 0x90; if (c == 0xF4) hi = 0x8F; }
  else return 0xFFFD;
  if (need >= n) return 0xFFFD;
  uint32_t c1 = s[1];
  if (c1 < lo || c1 > hi) return 0xFFFD;
  r = (r << 6) | (c1 & 0x3F);
  for (uint32_t k = 2; k <= need; k++) {
    uint32_t ck = s[k];
    if ((ck & 0xC0) != 0x80) return 0xFFFD;
    r = (r << 6) | (ck & 0x3F);
  }
  *width = need + 1;
  return r;
}

// unicode.IsSpace
M2K_HD bool is_space_rune(uint32_t r) {
  if (r < 0x80) return r == 0x20 || (r >= 0x09 && r <= 0x0D);
  if (r == 0x85 || r == 0xA0 || r == 0x1680 || r == 0x2028 || r == 0x2029 || r == 0x202F || r == 0x205F || r == 0x3000) return true;
  return r >= 0x2000 && r <= 0x200A;
}

// Representative of r's simple-case-folding class: strings.EqualFold(a, b) holds exactly when the
// two rune sequences map to the same representatives.  `from`/`to` is the generated table
// (casefold_table.h), sorted by `from`.
M2K_HD uint32_t fold_rune(uint32_t r, const uint32_t* from, const uint32_t* to, int n) {
  if (r < 0x80) return (r - 'A' < 26u) ? r + 32 : r;
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1;
    uint32_t v = from[mid];
    if (v == r) return to[mid];
    if (v < r) lo = mid + 1; else hi = mid - 1;
  }
  return r;
}

// 64-bit running hash over folded runes (FNV-1a style, one step per rune).
M2K_HD uint64_t hash_step(uint64_t h, uint32_t folded_rune) {
  h ^= folded_rune;
  h *= 0x100000001b3ull;
  return h;
}
M2K_HD uint64_t hash_finish(uint64_t h) {  // avalanche before table probing
  h ^= h >> 33; h *= 0xff51afd7ed558ccdull; h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ull; h ^= h >> 33;
  return h ? h : 1;
}

M2K_HD uint64_t hash_fold_bytes(uint64_t h, const uint8_t* s, size_t n, const uint32_t* from, const uint32_t* to, int nt) {
  size_t i = 0;
  while (i < n) {
    uint32_t w;
    uint32_t r = decode_rune(s + i, n - i, &w);
    h = hash_step(h, fold_rune(r, from, to, nt));
    i += w;
  }
  return h;
}

}  // namespace m2k
