// TEST INFRASTRUCTURE, never shipped: the device-side isDockerFile functions (move2kube_b200/csrc/dockerfile_dev.cuh --
// the very source nvcc compiles into k_df_head / k_df_states / k_df_validate) compiled for the HOST, so that the CPU
// tier can fuzz them against the oracle by the hundred thousand.  What is NOT covered here is what only exists on the
// device: how the kernels hand stripes to threads, the look-back between tiles and the byte scan's records; the GPU
// tier covers those.  The driver below does what the three kernels do, one stripe after the other:
//   the byte scan's first-word pre-filter -> head (k_df_head) -> composed line-state map per stripe of the body,
//   prefix composition of the maps (k_df_states; also folded in a second, tree-shaped order: the scan relies on composition being associative)
//   -> validation of every stripe from the state it starts in (k_df_validate).
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#define __device__
#define __forceinline__ inline
#define __ffsll __builtin_ffsll
#include "../../move2kube_b200/csrc/dockerfile_dev.cuh"

namespace {
constexpr uint32_t F_LONG = 1u << 1, F_FROM = 1u << 2, F_DF = 1u << 3;  // M2K_FILE_* (include/m2kdetect.h), checked by the test

struct Padded {  // file bases are 16-byte aligned and zero padded to 16 bytes
  uint8_t* p;
  Padded(const uint8_t* d, uint64_t n) {
    const size_t cap = (size_t)((n + 15) & ~(uint64_t)15) + 16;
    p = (uint8_t*)aligned_alloc(16, cap);
    memset(p, 0, cap);
    if (n) memcpy(p, d, n);
  }
  ~Padded() { free(p); }
};

bool long_line(const uint8_t* d, uint64_t n) {  // what the byte scan's records fold to (kernels.cu acc_long)
  uint64_t run = 0;
  for (uint64_t i = 0; i < n; i++) {
    if (d[i] == '\n') run = 0;
    else if (++run >= 65536) return true;
  }
  return false;
}

uint32_t tree_fold(const std::vector<uint32_t>& m, size_t lo, size_t hi) {  // composition of m[lo, hi) by halves
  if (hi - lo == 1) return m[lo];
  const size_t mid = lo + (hi - lo + 1) / 2;
  return m2k::df::map_compose(tree_fold(m, lo, mid), tree_fold(m, mid, hi));
}
}  // namespace

// flags of one file the way the device computes them, body cut into stripes of `stripe` bytes.
// Bit 31 set: the tree-shaped fold of the maps disagreed with the sequential one (must never happen).
extern "C" uint32_t dfhost_flags(const uint8_t* data, uint64_t len, uint32_t stripe) {
  using namespace m2k::df;
  if (stripe == 0) stripe = kStripe;
  uint32_t flags = 0;
  if (long_line(data, len)) return F_LONG;  // such files never reach k_df_head (kernels.cu finish_file)
  if (len == 0) return 0;
  Padded buf(data, len);
  {  // the byte scan's pre-filter on the file's first 16 bytes: a file it rules out never reaches k_df_head
    uint32_t w[4];
    memcpy(w, buf.p, 16);
    if (first_word_rules_out(w)) return flags;
  }
  uint64_t body = 0;
  uint32_t esc = '\\';
  if (!head(buf.p, len, false, &body, &esc)) return flags;
  flags |= F_FROM;
  const uint64_t n_stripes = (len - body + stripe - 1) / stripe;
  if (n_stripes == 0) return flags | F_DF;
  File f{buf.p, len, esc};
  std::vector<uint32_t> maps(n_stripes);
  std::vector<uint8_t> state(n_stripes);
  for (uint64_t k = 0; k < n_stripes; k++) {
    const uint64_t lo = body + k * stripe, hi = lo + stripe < len ? lo + stripe : len;
    maps[k] = stripe_map(f, body, lo, hi);
  }
  uint32_t s = 0;
  for (uint64_t k = 0; k < n_stripes; k++) { state[k] = (uint8_t)s; s = map_apply(maps[k], s); }
  if (map_apply(tree_fold(maps, 0, n_stripes), 0) != s) flags |= 1u << 31;
  bool ok = true;
  for (uint64_t k = 0; k < n_stripes && ok; k++) {
    const uint64_t lo = body + k * stripe, hi = lo + stripe < len ? lo + stripe : len;
    ok = stripe_validate(f, body, lo, hi, state[k]);
  }
  return ok ? flags | F_DF : flags;
}

// the head on a prefix [0, n_here) of the file, as a GPU that holds only the first part of a cut file runs it
// (k_df_head, PIECE_HEAD_*): returns 0 = verdict not final (rests on the last line present), 1 = FROM first,
// 2 = not FROM first.
extern "C" uint32_t dfhost_head_prefix(const uint8_t* data, uint64_t n_here) {
  Padded buf(data, n_here);
  uint64_t body = 0;
  uint32_t esc = '\\';
  bool touched_end = false;
  const bool from = m2k::df::head(buf.p, n_here, false, &body, &esc, &touched_end);
  return touched_end ? 0u : (from ? 1u : 2u);
}

extern "C" uint32_t dfhost_default_stripe(void) { return m2k::df::kStripe; }

// ---- the newline-run monoid (move2kube_b200/csrc/runacc.h) --------------------------------------------------------
#include "../../move2kube_b200/csrc/runacc.h"

namespace {
m2k::RunAcc span_summary(const uint8_t* d, uint64_t n, uint64_t limit) {  // straight from the bytes
  m2k::RunAcc r{n, n, n, 0};
  uint64_t first = n, last = n;
  for (uint64_t i = 0; i < n; i++) if (d[i] == '\n') { if (first == n) first = i; last = i; }
  if (first == n) return r;
  r.lead = first;
  r.trail = n - 1 - last;
  uint64_t run = 0;
  bool open = true;  // the run before the first newline is open on the left
  for (uint64_t i = 0; i < n; i++) {
    if (d[i] == '\n') { if (!open && run >= limit) r.longf = 1; run = 0; open = false; }
    else run++;
  }
  return r;
}
m2k::RunAcc fold_tree(const uint8_t* d, const uint64_t* cut, uint32_t lo, uint32_t hi, uint64_t limit, uint32_t* seed) {
  if (hi - lo == 1) return span_summary(d + cut[lo], cut[hi] - cut[lo], limit);
  *seed = *seed * 1664525u + 1013904223u;
  const uint32_t mid = lo + 1 + (*seed >> 8) % (hi - lo - 1);
  const m2k::RunAcc a = fold_tree(d, cut, lo, mid, limit, seed), b = fold_tree(d, cut, mid, hi, limit, seed);
  return m2k::acc_merge_at(a, b, limit);
}
}  // namespace

// data[0, n) cut at cut[0] = 0 < ... <= cut[n_cut] = n into spans; the spans' summaries merged in a random tree
// shape must equal the summary of the whole, and LONG_LINE must be "some line has >= limit bytes".
// Returns 0 when everything agrees, else a code saying what did not.
extern "C" uint32_t dfhost_runacc_check(const uint8_t* data, uint64_t n, const uint64_t* cut, uint32_t n_cut, uint64_t limit, uint32_t seed) {
  const m2k::RunAcc whole = span_summary(data, n, limit);
  const m2k::RunAcc got = n_cut ? fold_tree(data, cut, 0, n_cut, limit, &seed) : m2k::RunAcc{0, 0, 0, 0};
  if (n && (got.bytes != whole.bytes || got.lead != whole.lead || got.trail != whole.trail || got.longf != whole.longf)) return 1;
  uint64_t run = 0;
  uint32_t want = 0;
  for (uint64_t i = 0; i < n; i++) { if (data[i] == '\n') run = 0; else if (++run >= limit) want = 1; }
  if (m2k::acc_long_at(got.bytes, got.lead, got.trail, got.longf, limit) != want) return 2;
  return 0;
}
