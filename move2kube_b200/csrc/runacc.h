// Newline runs of a span of bytes, as a monoid: what the byte scan's 8-byte chunk records fold to per file
// (k_file_combine), per block share of a multi-GiB file, and across the pieces of a file cut between GPUs
// (k_merge_pieces).  bufio.Scanner fails on a line of >= 65536 bytes (bufio.MaxScanTokenSize; buildkit's
// parser.Parse hands that error on, internal/source/dockerfile2kube.go:124-128): LONG_LINE.
// Host and device: the CPU tier checks associativity and the answer on random spans (tests/test_df_host_fuzz.py).
#pragma once
#include <stdint.h>

#include "gostr.h"  // M2K_HD

namespace m2k {

struct RunAcc {            // monoid over a span of bytes of one file
  uint64_t bytes;          // bytes in the span
  uint64_t lead, trail;    // bytes before the first / after the last newline (== bytes when there is none)
  uint32_t longf;          // a line closed inside the span has >= the limit
};

// a then b; `limit` = kLongLine in the product (a parameter so that the property test can use short lines)
M2K_HD RunAcc acc_merge_at(const RunAcc& a, const RunAcc& b, uint64_t limit) {
  if (a.bytes == 0) return b;
  if (b.bytes == 0) return a;
  RunAcc r;
  r.bytes = a.bytes + b.bytes;
  r.longf = a.longf | b.longf;
  const bool a_free = a.lead == a.bytes, b_free = b.lead == b.bytes;
  r.lead = a_free ? a.bytes + b.lead : a.lead;
  r.trail = b_free ? b.bytes + a.trail : b.trail;
  if (!a_free && !b_free && a.trail + b.lead >= limit) r.longf = 1;  // a line closed on both sides
  return r;
}

// LONG_LINE of a whole file: a closed line, or the open runs at the file's start / end
M2K_HD uint32_t acc_long_at(uint64_t bytes, uint64_t lead, uint64_t trail, uint32_t longf, uint64_t limit) {
  if (longf) return 1;
  if (lead >= limit) return 1;  // also the file without any newline (lead == bytes)
  return (lead != bytes && trail >= limit) ? 1u : 0u;
}

}  // namespace m2k
