// Procedural content of the synthetic trees.
// Bytes are a pure function of (seed, file id, file class, length, inject position, offset), so the
// host fill (m2k_synth_fill_host, read by the CPU oracle) and the device fill
// (m2k_synth_fill_device) produce identical content without 100 GB ever touching a disk.
#pragma once
#include <stdint.h>
#include "gostr.h"

namespace m2k {

enum SynthClass : uint8_t {
  SC_TEXT = 0,        // printable lines, one '\n' per 64 bytes on average
  SC_BINARY = 1,      // raw random bytes
  SC_NONL = 2,        // text without any newline (minified): a >= 64 KiB line when long enough
  SC_DOCKERFILE = 3,  // valid Dockerfile head, text body
  SC_NEAR_ARGS = 4,   // only ARG lines before text: first instruction is not FROM
  SC_NEAR_FROMX = 5,  // "FROMX base"
  SC_NEAR_TWOARGS = 6,// "FROM base extra": instruction is FROM but the regexp rejects it
  SC_NEAR_ENVERR = 7  // "FROM base" then "ENV BROKEN": parse error later in the file
};

M2K_HD uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

M2K_HD uint64_t synth_file_key(uint64_t seed, uint32_t file_id) {
  return splitmix64(seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(file_id + 1)));
}

M2K_HD uint32_t synth_char(uint32_t v) {  // 64-symbol alphabet
  if (v < 26) return 'a' + v;
  if (v < 52) return 'A' + (v - 26);
  if (v < 62) return '0' + (v - 52);
  return v == 62 ? ' ' : '_';
}

// 64-byte heads (exactly 64 characters each, newline terminated)
M2K_HD const char* synth_head(uint8_t cls) {
  switch (cls) {
    case SC_DOCKERFILE:   return "FROM registry.example.com/base/image:1.0 AS build\nRUN make all \n";
    case SC_NEAR_ARGS:    return "ARG VERSION=1.0\nARG TARGET\n# only build arguments in this file \n";
    case SC_NEAR_FROMX:   return "FROMX registry.example.com/base/image:1.0\nRUN make all  install\n";
    case SC_NEAR_TWOARGS: return "FROM registry.example.com/base/image extra\nRUN make all install\n";
    case SC_NEAR_ENVERR:  return "FROM registry.example.com/base/image:1.0\nENV BROKEN\nRUN make al\n";
    default: return nullptr;
  }
}

// 16 bytes of file `file_id` at block index b (bytes [16b, 16b+16)); zero beyond `len`.
M2K_HD void synth_block16(uint64_t key, uint8_t cls, uint64_t len, uint64_t inject_pos, uint64_t b, uint8_t out[16]) {
  uint64_t base = b * 16;
  uint64_t r0 = splitmix64(key + 2 * b), r1 = splitmix64(key + 2 * b + 1);
  if (cls == SC_BINARY) {
    for (int i = 0; i < 8; i++) { out[i] = (uint8_t)(r0 >> (8 * i)); out[8 + i] = (uint8_t)(r1 >> (8 * i)); }
  } else {
    for (int i = 0; i < 10; i++) out[i] = (uint8_t)synth_char((uint32_t)(r0 >> (6 * i)) & 63);
    for (int i = 0; i < 6; i++) out[10 + i] = (uint8_t)synth_char((uint32_t)(r1 >> (6 * i)) & 63);
    if (cls != SC_NONL && ((r1 >> 40) & 3) == 0) out[(r1 >> 44) & 15] = '\n';
    const char* head = synth_head(cls);
    if (head && base < 64) for (int i = 0; i < 16; i++) out[i] = (uint8_t)head[base + i];
  }
  if (inject_pos != ~0ull && inject_pos < base + 16 && inject_pos + 8 > base) {
    const char* m = "__main__";
    for (int i = 0; i < 16; i++) {
      uint64_t p = base + i;
      if (p >= inject_pos && p < inject_pos + 8) out[i] = (uint8_t)m[p - inject_pos];
    }
  }
  if (base + 16 > len) for (int i = 0; i < 16; i++) if (base + i >= len) out[i] = 0;
}

}  // namespace m2k
