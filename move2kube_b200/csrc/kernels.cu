// sm_100a kernels of the source-detection path.  HBM-bound integer/byte work: the rules that
// matter are coalesced 16-byte loads, enough bytes in flight per SM, and grids sized in multiples
// of the SM count (DESIGN.md has the per-kernel rooflines).
#include <cuda_runtime.h>
#include <stdint.h>

#include "casefold_table.h"
#include "dockerfile_dev.cuh"
#include "gostr.h"
#include "m2k_internal.h"
#include "runacc.h"
#include "synth.h"

namespace m2k {

static_assert(df::kStripe == kDfStripe, "stripe size");

#define FULL 0xFFFFFFFFu

// A checked build (-DM2K_CHECKED=1, tools/scan_variants.py "checked") counts every access that falls outside
// the buffer it belongs to instead of making it: our own bounds checks, run over the adversarial parity tests,
// where a memory checker cannot be used.  Slots: 0 scan load outside the shard's bytes, 1 chunk record index,
// 2 head list / candidate index, 3 stripe or tile index, 4 a Dockerfile stripe outside its file.
#ifndef M2K_CHECKED
#define M2K_CHECKED 0
#endif
__device__ unsigned long long g_violations[8];
#if M2K_CHECKED
#define M2K_CHECK(cond, slot) ((cond) ? true : (atomicAdd(&g_violations[slot], 1ull), false))
#else
#define M2K_CHECK(cond, slot) true
#endif
int read_violations(unsigned long long* out) {  // host: 0 or a cudaError_t; zeroes the counters
  unsigned long long zero[8] = {};
  cudaError_t e = cudaMemcpyFromSymbol(out, g_violations, sizeof zero);
  if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_violations, zero, sizeof zero);
  return (int)e;
}
int checked_build() { return M2K_CHECKED; }

// ---------------------------------------------------------------------------------------------
// small helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

__global__ void k_fill_u32(uint32_t* p, uint32_t v, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = v;
}

void launch_fill_u32(Ctx* c, uint32_t* p, uint32_t v, size_t n) {
  if (!n) return;
  int blocks = (int)((n + 255) / 256);
  if (blocks > c->sm_count * 8) blocks = c->sm_count * 8;
  k_fill_u32<<<blocks, 256, 0, c->stream>>>(p, v, n);
  c->launches++;
}

// ---------------------------------------------------------------------------------------------
// Kernel 1: name features.  One thread per directory entry; the signature table (exact names,
// suffixes) is staged in shared memory; a warp-level segmented OR (match_any + redux) folds the
// entries of one directory before a single atomicOr per (warp, directory) -- and the same lane then
// pushes the bits up the directory's ancestor chain (dir_subtree = OR over the subtree: what the
// scripts' `find "$1"/. -name` computes by re-walking every subtree).
// Replaces: [ -f D/name ] / find -name "*.ext" / "$1/"*.war of the 18 detect scripts
// (internal/assets/*/m2k*detect.sh) and GetFilesByName's basename test (utils.go:108-113).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_name_features(
    uint32_t n_entries, const uint32_t* __restrict__ entry_dir, const uint32_t* __restrict__ entry_name_off,
    const uint8_t* __restrict__ entry_type, const uint8_t* __restrict__ names, const NameSig* __restrict__ sigs, int n_sigs,
    const uint32_t* __restrict__ dir_parent, uint32_t* __restrict__ entry_feat, uint32_t* __restrict__ dir_direct,
    uint32_t* __restrict__ dir_subtree, uint32_t* __restrict__ dir_first_war, uint32_t* __restrict__ special, uint32_t special_cap) {
  __shared__ NameSig s_sigs[kMaxNameSigs];
  for (int i = threadIdx.x; i < n_sigs; i += blockDim.x) s_sigs[i] = sigs[i];
  __syncthreads();
  uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  bool valid = e < n_entries;
  uint32_t feat = 0, dir = M2K_NONE;
  if (valid) {
    dir = entry_dir[e];
    uint32_t off = entry_name_off[e], len = entry_name_off[e + 1] - off;
    uint32_t type = entry_type[e];
    uint32_t lt = type & 0xF, ft = type >> 4;
    // the last <=16 bytes of the name, right aligned: byte 0 of lo is the last character
    uint64_t lo = 0, hi = 0;
    uint32_t take = len < 16 ? len : 16;
    for (uint32_t i = 0; i < take; i++) {
      uint64_t ch = names[off + len - 1 - i];
      if (i < 8) lo |= ch << (8 * i); else hi |= ch << (8 * (i - 8));
    }
    uint32_t first = len ? names[off] : 0;
    if (lt != M2K_T_ERR) {
      for (int s = 0; s < n_sigs; s++) {
        const NameSig& g = s_sigs[s];
        bool m = (g.kind == 0) ? (len == g.len) : (len >= g.len);
        m = m && ((lo & g.mask_lo) == g.lo) && ((hi & g.mask_hi) == g.hi);
        if (!m) continue;
        if (g.need == 1 && ft != M2K_T_REG) continue;
        if (g.need == 2 && lt == M2K_T_DIR) continue;
        if (g.need == 3 && first == '.') continue;
        feat |= 1u << g.feature;
      }
    }
    entry_feat[e] = feat;
    if (feat & ((1u << M2K_N_IGNOREFILE) | (1u << M2K_N_USER_DF) | (1u << M2K_N_USER_S2I))) {
      // rare: list these entries so the host need not look through every feature word
      uint32_t k = atomicAdd(&special[0], 1u);
      if (k < special_cap) special[1 + k] = e;
    }
    if (feat & (1u << M2K_N_WAR)) atomicMin(&dir_first_war[dir], e);  // entries are name-sorted inside a directory
  }
  // segmented OR over the lanes that share a directory
  uint32_t peers = __match_any_sync(FULL, dir);
  uint32_t dirfeat = feat & ~(1u << M2K_N_WAR);  // the war bit of a directory is decided from dir_first_war
  uint32_t ored = __reduce_or_sync(peers, dirfeat);
  if (valid && ored && (uint32_t)(__ffs(peers) - 1) == (threadIdx.x & 31)) {
    atomicOr(&dir_direct[dir], ored);
    // find's recursion, once for all directories: the bits climb the ancestor chain.  A level that already
    // holds all of them ends the climb -- whoever put a bit there first carries it further up (it saw the
    // bit missing, so it went on), which makes the whole pass close to linear in the number of entries.
    for (uint32_t x = dir;;) {
      const uint32_t old = atomicOr(&dir_subtree[x], ored);
      if ((old & ored) == ored || x == 0) break;
      x = dir_parent[x];
    }
  }
}

void launch_name_features(Ctx* c, DevTree* t) {
  if (!t->n_entries) return;
  uint32_t blocks = (t->n_entries + 255) / 256;
  k_name_features<<<blocks, 256, 0, c->stream>>>(t->n_entries, t->entry_dir, t->entry_name_off, t->entry_type, t->names,
                                                 c->d_sigs, c->n_sigs, t->dir_parent, t->entry_feat, t->dir_direct, t->dir_subtree,
                                                 t->dir_first_war, t->special, t->special_cap);
  c->launches++;
}

// ---------------------------------------------------------------------------------------------
// Kernel 2: byte scan -- the HBM-roofline kernel.  One warp per 64 KiB chunk of one file; a warp
// row is 32 lanes x 16 bytes = 512 contiguous bytes, 8 rows in flight per lane.
//   * "__main__" (python detect scripts :21/:22): any occurrence fully contains one 4-byte ALIGNED
//     word equal to "__ma", "_mai", "main" or "ain_", so the streaming test is four compares per
//     word; the rare hit is confirmed bytewise from global memory (file bounds respected).
//   * line length (bufio.Scanner's 64 KiB cap inside dockerparser.Parse, dockerfile2kube.go:124): per
//     16-byte unit "has a newline" (SWAR), ballot per row; only the chunk's first and last newline
//     matter, and their positions are taken to the byte by the one lane that holds them.  What leaves
//     the chunk is an 8-byte record (bytes before the first / after the last newline, flag bits):
//     exact, so k_file_combine -- or another GPU, for a file cut between shards -- stitches records
//     without ever reading the file again.
// Algorithmic traffic: every content byte read once + 8 B written per 64 KiB.
// ---------------------------------------------------------------------------------------------
// tunables (tools/scan_variants.py builds and times alternatives; these defaults won on B200)
#ifndef M2K_SCAN_ROWS
#define M2K_SCAN_ROWS 8
#endif
#ifndef M2K_SCAN_THREADS
#define M2K_SCAN_THREADS 256
#endif
#ifndef M2K_SCAN_MINBLOCKS
#define M2K_SCAN_MINBLOCKS 4
#endif
#ifndef M2K_SCAN_WAVES
#define M2K_SCAN_WAVES 32
#endif
#ifndef M2K_SCAN_PREFETCH
#define M2K_SCAN_PREFETCH 0
#endif
#ifndef M2K_SCAN_QUOTA_WAVES
#define M2K_SCAN_QUOTA_WAVES 0  // ticket mode: 0 = one wave of blocks that never retire; W = W waves, warps retire after a quota
#endif
#ifndef M2K_SCAN_TICKET
#define M2K_SCAN_TICKET 1  // one resident wave of blocks, further chunks drawn from a counter (13.60 vs 13.75 ms on config C, 0.163 vs 0.171 on B)
#endif
constexpr int kScanRows = M2K_SCAN_ROWS;        // rows (512 B each) loaded before use
constexpr int kScanThreads = M2K_SCAN_THREADS;

__device__ __forceinline__ uint32_t has_newline(uint4 v) {
  const uint32_t nl = 0x0A0A0A0Au, ones = 0x01010101u;
  uint32_t x0 = v.x ^ nl, x1 = v.y ^ nl, x2 = v.z ^ nl, x3 = v.w ^ nl;
  uint32_t t = ((x0 - ones) & ~x0) | ((x1 - ones) & ~x1) | ((x2 - ones) & ~x2) | ((x3 - ones) & ~x3);
  return t & 0x80808080u;
}

// exact per-byte test: bit 7 of every byte of w that is '\n'
__device__ __forceinline__ uint32_t nl_bits(uint32_t w) {
  uint32_t x = w ^ 0x0A0A0A0Au;
  return ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
}
// byte index (0..15) of the first / last newline of a unit that has one
__device__ __noinline__ uint32_t first_nl_byte(uint4 v) {
  uint32_t m;
  if ((m = nl_bits(v.x))) return (uint32_t)(__ffs(m) - 1) >> 3;
  if ((m = nl_bits(v.y))) return 4 + ((uint32_t)(__ffs(m) - 1) >> 3);
  if ((m = nl_bits(v.z))) return 8 + ((uint32_t)(__ffs(m) - 1) >> 3);
  m = nl_bits(v.w);
  return 12 + ((uint32_t)(__ffs(m) - 1) >> 3);
}
__device__ __noinline__ uint32_t last_nl_byte(uint4 v) {
  uint32_t m;
  if ((m = nl_bits(v.w))) return 12 + ((31u - (uint32_t)__clz(m)) >> 3);
  if ((m = nl_bits(v.z))) return 8 + ((31u - (uint32_t)__clz(m)) >> 3);
  if ((m = nl_bits(v.y))) return 4 + ((31u - (uint32_t)__clz(m)) >> 3);
  m = nl_bits(v.x);
  return (31u - (uint32_t)__clz(m)) >> 3;
}

__device__ __forceinline__ bool main_word(uint32_t w) {
  // little-endian words of "__ma" "_mai" "main" "ain_"
  return w == 0x616D5F5Fu || w == 0x69616D5Fu || w == 0x6E69616Du || w == 0x5F6E6961u;
}

// exact check around an aligned word hit at file offset `pos`
__device__ __noinline__ bool confirm_main(const uint8_t* file, uint64_t len, uint64_t pos) {
  const char* pat = "__main__";
  for (int back = 0; back <= 3; back++) {
    if (pos < (uint64_t)back) break;
    uint64_t s = pos - back;
    if (s + 8 > len) continue;
    bool ok = true;
    for (int i = 0; i < 8; i++) ok = ok && file[s + i] == (uint8_t)pat[i];
    if (ok) return true;
  }
  return false;
}

// First 16 bytes of a file (lane 0 of chunk 0): can the first instruction still be FROM?  Returns
// true when it cannot -- the first word, read past blank lines, is neither a keyword of the
// Dockerfile head (from / arg / env / label / onbuild) nor possibly continued -- so isDockerFile is
// false without ever touching the file again.  Anything unusual (comment or directive, non-ASCII
// byte, escape character, NUL, no verdict within 16 bytes) is left to k_df_head.
__device__ __noinline__ bool first_word_rules_out_dockerfile(uint4 v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  return df::first_word_rules_out(w);  // dockerfile_dev.cuh, where the host-side fuzz of the parsing code reaches it
}

struct ScanArgs {
  uint32_t n_chunks, n_shard_files;
  const uint32_t* chunk_first; const uint64_t* file_off; const uint64_t* file_len;
  uint32_t file_begin;
  const uint8_t* bytes; uint64_t shard_byte0;
  uint32_t first_skip;
  ChunkRec* chunk_rec;
  uint32_t* ticket;  // work items are drawn from this counter (NULL: strided over the grid)
  uint32_t quota;    // ticket mode: a warp retires after this many chunks (0: never) -- see launch_scan_bytes
  uint64_t shard_bytes;  // resident bytes (checked builds)
  const uint32_t* chunk_file;  // [n_chunks] shard-local index of the file each chunk belongs to (NULL: bisect chunk_first)
};
__device__ __forceinline__ bool in_shard(const ScanArgs& a, const void* p, size_t n) {
  const uint64_t off = (uint64_t)((const uint8_t*)p - a.bytes);  // wraps to a huge value below the buffer
  return off <= a.shard_bytes && n <= a.shard_bytes - off;
}

// where work item `item` (a 64 KiB chunk of one file) lies
struct ChunkCtx {
  const uint8_t* fbase;  // byte 0 of the file; may point before the shard, only [start - 16, ...) is touched
  uint64_t flen, start;  // file length; offset of the chunk in the file
  uint32_t range, n_units;
  const uint4* src;
};
__device__ __forceinline__ ChunkCtx locate_chunk(const ScanArgs& a, uint32_t item) {
  // which file: last f with chunk_first[f] <= item (empty files share their successor's value).  A table made
  // at upload answers in one load; without it (the no-allocation entry point) twenty dependent loads bisect
  // chunk_first -- for a tree of tiny files that search cost more than the bytes.
  uint32_t lo = 0;
  if (a.chunk_file) {
    lo = a.chunk_file[item];
  } else {
    uint32_t hi = a.n_shard_files;
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (a.chunk_first[mid] <= item) lo = mid; else hi = mid; }
  }
  const uint32_t f = a.file_begin + lo;
  ChunkCtx c;
  c.flen = a.file_len[f];
  // the shard may begin `first_skip` chunks into its first file (the rest of it is on another GPU)
  c.start = (uint64_t)(item - a.chunk_first[lo] + (lo == 0 ? a.first_skip : 0u)) * kChunkBytes;
  c.fbase = a.bytes + (a.file_off[f] - a.shard_byte0);
  c.range = (uint32_t)(c.flen - c.start < kChunkBytes ? c.flen - c.start : kChunkBytes);
  c.n_units = (c.range + 15) >> 4;
  c.src = reinterpret_cast<const uint4*>(c.fbase + c.start);
  return c;
}

// Inside one 64 KiB chunk a newline-free run that is closed on both sides is shorter than 64 KiB,
// so only the first and the last newline of the chunk matter: the first is looked for until it is
// found, the last in the final batch of rows (walking back from there in the rare case it has none).
struct ChunkAcc { uint32_t lead, last_nl, found, not_df; };  // byte offsets inside the chunk; M2K_NONE: not seen

// one batch of kScanRows rows (row r, lane l holds unit u0 + r * 32 + l; units past the end are zero)
__device__ __forceinline__ void scan_batch(const uint4 (&v)[kScanRows], uint32_t u0, uint32_t lane, const ChunkCtx& c, ChunkAcc& acc) {
  if (c.start == 0 && u0 == 0 && lane == 0) acc.not_df = first_word_rules_out_dockerfile(v[0]) ? 1u : 0u;
  bool hit_any = false;
#pragma unroll
  for (int r = 0; r < kScanRows; r++)
    hit_any = hit_any || main_word(v[r].x) || main_word(v[r].y) || main_word(v[r].z) || main_word(v[r].w);
  if (__any_sync(FULL, hit_any)) {  // rare: confirm bytewise, file bounds respected
    bool ok = false;
#pragma unroll
    for (int r = 0; r < kScanRows; r++) {
      uint64_t pos = c.start + (uint64_t)(u0 + r * 32 + lane) * 16;
      if (main_word(v[r].x)) ok = ok || confirm_main(c.fbase, c.flen, pos);
      if (main_word(v[r].y)) ok = ok || confirm_main(c.fbase, c.flen, pos + 4);
      if (main_word(v[r].z)) ok = ok || confirm_main(c.fbase, c.flen, pos + 8);
      if (main_word(v[r].w)) ok = ok || confirm_main(c.fbase, c.flen, pos + 12);
    }
    if (__any_sync(FULL, ok)) acc.found = 1;
  }
  if (acc.lead == M2K_NONE) {  // warp-uniform: first newline, rows in order
#pragma unroll
    for (int r = 0; r < kScanRows; r++) {
      if (acc.lead == M2K_NONE) {
        uint32_t B = __ballot_sync(FULL, has_newline(v[r]) != 0);  // rows past the end are zero: no newline
        if (B) {
          int l = __ffs(B) - 1;
          acc.lead = (u0 + r * 32 + l) * 16 + __shfl_sync(FULL, first_nl_byte(v[r]), l);
        }
      }
    }
  }
  if (u0 + 32 * kScanRows >= c.n_units) {  // warp-uniform: last newline, rows in reverse
#pragma unroll
    for (int r = kScanRows - 1; r >= 0; r--) {
      if (acc.last_nl == M2K_NONE) {
        uint32_t B = __ballot_sync(FULL, has_newline(v[r]) != 0);
        if (B) {
          int l = 31 - __clz(B);
          acc.last_nl = (u0 + r * 32 + l) * 16 + __shfl_sync(FULL, last_nl_byte(v[r]), l);
        }
      }
    }
  }
}

// the chunk's 8-byte record, once all its batches are through
__device__ __forceinline__ ChunkRec chunk_record(const ChunkCtx& c, ChunkAcc acc, uint32_t lane) {
  uint32_t trail;
  if (acc.lead == M2K_NONE) {
    acc.lead = trail = c.range;  // no newline in the whole chunk
  } else {
    if (acc.last_nl == M2K_NONE) {
      // the last batch has no newline (rare): walk back row by row; the newline at `lead` stops the search
      uint32_t ubase = ((c.n_units - 1) / (32 * kScanRows)) * (32 * kScanRows);
      while (acc.last_nl == M2K_NONE) {
        ubase -= 32;
        uint4 w = ldg_stream(c.src + ubase + lane);
        uint32_t B = __ballot_sync(FULL, has_newline(w) != 0);
        if (B) {
          int l = 31 - __clz(B);
          acc.last_nl = (ubase + l) * 16 + __shfl_sync(FULL, last_nl_byte(w), l);
        }
      }
    }
    trail = c.range - 1 - acc.last_nl;
  }
  return rec_pack(acc.lead, trail, acc.found, acc.not_df);
}

__global__ void __launch_bounds__(kScanThreads, M2K_SCAN_MINBLOCKS) k_scan_bytes(ScanArgs a) {
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t warps_per_block = kScanThreads / 32;
  const uint32_t n_warps = gridDim.x * warps_per_block;
  uint32_t item = blockIdx.x * warps_per_block + (threadIdx.x >> 5), taken = 0;
  if (a.ticket) {
    uint32_t nx = 0;
    if (lane == 0) nx = atomicAdd(a.ticket, 1u);
    item = __shfl_sync(FULL, nx, 0);
  }
  while (item < a.n_chunks) {
    const ChunkCtx c = locate_chunk(a, item);
    ChunkAcc acc{M2K_NONE, M2K_NONE, 0, 0};
    for (uint32_t u0 = 0; u0 < c.n_units; u0 += 32 * kScanRows) {
      uint4 v[kScanRows];
#pragma unroll
      for (int r = 0; r < kScanRows; r++) {
        uint32_t u = u0 + r * 32 + lane;
        v[r] = (u < c.n_units && M2K_CHECK(in_shard(a, c.src + u, 16), 0)) ? ldg_stream(c.src + u) : make_uint4(0, 0, 0, 0);
      }
#if M2K_SCAN_PREFETCH
      // the next batch on its way into L2 while this one is looked at (costs no registers): one request per
      // 128-byte line, from the lane that holds its first 16 bytes
      if ((lane & 7u) == 0) {
#pragma unroll
        for (int r = 0; r < kScanRows; r++) {
          uint32_t u = u0 + (kScanRows + r) * 32 + lane;
          if (u < c.n_units) asm volatile("prefetch.global.L2 [%0];" ::"l"(c.src + u));
        }
      }
#endif
      scan_batch(v, u0, lane, c, acc);
    }
    const ChunkRec rec = chunk_record(c, acc, lane);
    if (lane == 0 && M2K_CHECK(item < a.n_chunks, 1)) a.chunk_rec[item] = rec;
    if (a.ticket) {  // chunks differ in size (the last chunk of a file): handed out as warps come free
      if (a.quota && ++taken >= a.quota) break;
      uint32_t nx = 0;
      if (lane == 0) nx = atomicAdd(a.ticket, 1u);
      item = __shfl_sync(FULL, nx, 0);
    } else {
      item += n_warps;
    }
  }
}

// Measured alternative (tools/scan_variants.py, -DM2K_SCAN_BULK=1): the same scan fed by the bulk-copy
// engine.  Every warp owns a ring of kBulkStages stages of one batch (kScanRows x 512 bytes) in shared
// memory; lane 0 issues `cp.async.bulk` global -> shared copies that complete on the stage's mbarrier,
// kBulkStages batches ahead of where the warp is reading, across chunk boundaries (the warp's
// sequence of chunks is fixed: item, item + n_warps, ...).  Rows are then read from shared memory with
// conflict-free 16-byte loads.
#ifndef M2K_BULK_WARPS
#define M2K_BULK_WARPS 16
#endif
#ifndef M2K_BULK_STAGES
#define M2K_BULK_STAGES 3
#endif
constexpr uint32_t kBulkWarps = M2K_BULK_WARPS, kBulkStages = M2K_BULK_STAGES;
constexpr uint32_t kBulkStageBytes = kScanRows * 512;
constexpr size_t kBulkSmem = (size_t)kBulkWarps * kBulkStages * (kBulkStageBytes + 8);
static_assert(kBulkSmem <= 227 * 1024, "the rings of one block must fit the SM's shared memory");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n .reg .pred p;\n WAIT_%=:\n"
      " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      " @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// a cursor over the batches of a warp's chunk sequence
struct BatchCursor { uint32_t item, u0; ChunkCtx c; };

__global__ void __launch_bounds__(kBulkWarps * 32, 1) k_scan_bytes_bulk(ScanArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t n_warps = gridDim.x * kBulkWarps;
  uint8_t* ring = smem + (size_t)warp * kBulkStages * kBulkStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)kBulkWarps * kBulkStages * kBulkStageBytes) + warp * kBulkStages;
  if (lane == 0) for (uint32_t st = 0; st < kBulkStages; st++) mbar_init(&bars[st], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  const uint32_t first = blockIdx.x * kBulkWarps + warp;
  if (first >= a.n_chunks) return;
  // the fetch cursor runs kBulkStages batches ahead of the read cursor
  BatchCursor fc{first, 0, locate_chunk(a, first)};
  auto issue = [&](uint32_t stage) {  // copy the fetch cursor's batch into `stage`, advance the cursor
    if (fc.item >= a.n_chunks) return;
    if (fc.c.n_units == 0) {  // cannot happen (an empty file has no chunk), kept for safety
      fc.item = a.n_chunks; return;
    }
    const uint32_t units = fc.c.n_units - fc.u0 < 32u * kScanRows ? fc.c.n_units - fc.u0 : 32u * kScanRows;
    if (lane == 0) bulk_load(ring + (size_t)stage * kBulkStageBytes, fc.c.src + fc.u0, units * 16, &bars[stage]);
    fc.u0 += 32 * kScanRows;
    if (fc.u0 >= fc.c.n_units) {
      fc.item += n_warps; fc.u0 = 0;
      if (fc.item < a.n_chunks) fc.c = locate_chunk(a, fc.item);
    }
  };
  for (uint32_t st = 0; st < kBulkStages; st++) issue(st);
  uint32_t stage = 0, parity = 0;
  for (uint32_t item = first; item < a.n_chunks; item += n_warps) {
    const ChunkCtx c = locate_chunk(a, item);
    ChunkAcc acc{M2K_NONE, M2K_NONE, 0, 0};
    for (uint32_t u0 = 0; u0 < c.n_units; u0 += 32 * kScanRows) {
      mbar_wait(&bars[stage], parity);
      const uint4* rows = reinterpret_cast<const uint4*>(ring + (size_t)stage * kBulkStageBytes);
      uint4 v[kScanRows];
#pragma unroll
      for (int r = 0; r < kScanRows; r++) {
        uint32_t u = u0 + r * 32 + lane;
        v[r] = u < c.n_units ? rows[r * 32 + lane] : make_uint4(0, 0, 0, 0);
      }
      __syncwarp();  // every lane has its rows in registers: the stage can be refilled
      issue(stage);
      if (++stage == kBulkStages) { stage = 0; parity ^= 1; }
      scan_batch(v, u0, lane, c, acc);
    }
    const ChunkRec rec = chunk_record(c, acc, lane);
    if (lane == 0) a.chunk_rec[item] = rec;
  }
}

void launch_scan_bytes(Ctx* c, DevTree* t, ChunkRec* chunk_rec, uint32_t* ticket) {
  if (!t->n_chunks) return;
  ScanArgs a{t->n_chunks, t->file_end - t->file_begin, t->chunk_first, t->file_off, t->file_len, t->file_begin,
             t->bytes, t->shard_byte0, t->first_skip, chunk_rec, ticket, 0, t->shard_bytes, t->chunk_file};
#if defined(M2K_SCAN_BULK) && M2K_SCAN_BULK
  static bool attr_set = false;
  if (!attr_set) { cudaFuncSetAttribute(k_scan_bytes_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBulkSmem); attr_set = true; }
  uint32_t want = (t->n_chunks + kBulkWarps - 1) / kBulkWarps;
  uint32_t cap = (uint32_t)c->sm_count;  // one block per SM, persistent over its warps' chunk sequences
  a.ticket = nullptr;
  k_scan_bytes_bulk<<<want < cap ? want : cap, kBulkWarps * 32, kBulkSmem, c->stream>>>(a);
#else
  uint32_t warps_per_block = kScanThreads / 32;
  uint32_t want = (t->n_chunks + warps_per_block - 1) / warps_per_block;
  uint32_t cap = (uint32_t)c->sm_count * M2K_SCAN_MINBLOCKS * M2K_SCAN_WAVES;  // resident blocks x waves
#if M2K_SCAN_TICKET
  if (a.ticket) {
#if M2K_SCAN_QUOTA_WAVES
    // Chunks come from a counter, but a warp retires after `quota` of them and the grid holds several
    // waves of blocks: a block lives for a fraction of the scan, so the high-priority blocks of the
    // directory-level kernels (side stream) find a free slot within that fraction instead of after the
    // whole scan -- which is what a single wave of never-retiring blocks makes them do.  The quota is
    // twice the even share: late blocks find the counter exhausted and leave at once.
    cap = (uint32_t)c->sm_count * M2K_SCAN_MINBLOCKS * M2K_SCAN_QUOTA_WAVES;
    const uint32_t grid_warps = (want < cap ? want : cap) * warps_per_block;
    a.quota = 2 * ((t->n_chunks + grid_warps - 1) / grid_warps);
#else
    cap = (uint32_t)c->sm_count * M2K_SCAN_MINBLOCKS;  // one resident wave, never retiring
#endif
  }
#else
  a.ticket = nullptr;
#endif
  uint32_t blocks = want < cap ? want : cap;
  k_scan_bytes<<<blocks, kScanThreads, 0, c->stream>>>(a);
#endif
  c->launches++;
}

// ---------------------------------------------------------------------------------------------
// Kernel 3: per-file combine of the chunk records (one thread per file; the whole warp for files of
// more than 128 chunks).  Produces HAS_MAIN and
// LONG_LINE, and the per-directory content partials for "__main__" in a file named *py.
// ---------------------------------------------------------------------------------------------
// RunAcc, acc_merge, acc_long: runacc.h (shared with the host-side property test of the fold)
__device__ __forceinline__ RunAcc acc_merge(const RunAcc& a, const RunAcc& b) { return acc_merge_at(a, b, (uint64_t)kLongLine); }
__host__ __device__ __forceinline__ uint32_t acc_long(uint64_t bytes, uint64_t lead, uint64_t trail, uint32_t longf) {
  return acc_long_at(bytes, lead, trail, longf, (uint64_t)kLongLine);
}

constexpr uint32_t kCombineWarpChunks = 128;  // files with more chunks are folded by the whole warp

// everything the per-file kernels (combine, Dockerfile heads, body stripes) look at
struct FileArgs {
  uint32_t n_shard_files, file_begin, first_skip, last_end;  // chunks of the first file before the shard; end chunk of the last file (M2K_NONE: all)
  const uint32_t* chunk_first; const ChunkRec* chunk_rec;
  const uint64_t* file_off; const uint64_t* file_len; const uint8_t* bytes; uint64_t shard_byte0;
  const uint32_t* file_entry; const uint32_t* entry_dir; const uint32_t* entry_name_off; const uint8_t* names;
  uint8_t* file_flags; uint32_t* dir_content; uint32_t* dir_main;
  PieceSummary* pieces;  // [2]: the shard's first / last file when only part of it is here
  // isDockerFile scratch
  DfCounters* ctr;
  uint32_t* head_list;   // shard-local indices of the files whose head must be parsed
  uint32_t* cand_file; uint64_t* cand_body; uint8_t* cand_esc; uint32_t* cand_stripe0; uint32_t* cand_err; uint32_t* cand_done;
  uint8_t* stripe_state; // per stripe (kDfStripe bytes) of candidate body: the line state at its first line start
  uint32_t* tile_state;  // per tile of 256 stripes: (pass number << 3) | 4: end state known | map (2 bits) or end state (bit 0)
  uint32_t epoch;
  uint32_t stripe_cap, tile_cap;  // sizes of the two tables (checked builds)
};

// fold records [k0, k1) (indices into the file's records present in this shard, the first of which
// covers file bytes from `byte0`); *has_main / *not_df accumulate the flag bits
__device__ __forceinline__ RunAcc fold_records(const FileArgs& a, uint32_t c0, uint32_t k0, uint32_t k1, uint64_t byte0, uint64_t len,
                                               uint32_t* has_main, uint32_t* not_df) {
  RunAcc acc{0, 0, 0, 0};
  auto take = [&](uint32_t k, ChunkRec rec) {
    uint64_t bstart = byte0 + (uint64_t)k * kChunkBytes;
    uint64_t nb = len - bstart < kChunkBytes ? len - bstart : kChunkBytes;
    RunAcc b{nb, rec_lead(rec), rec_trail(rec), 0};
    *has_main |= rec_main(rec);
    *not_df |= rec_not_df(rec);
    acc = acc_merge(acc, b);
  };
  uint32_t k = k0;
  // long runs (the share of a multi-GiB file's records one thread folds): eight loads in flight before
  // the merges, which depend on each other
  for (; k + 8 <= k1; k += 8) {
    ChunkRec r[8];
#pragma unroll
    for (int j = 0; j < 8; j++) r[j] = a.chunk_rec[c0 + k + j];
#pragma unroll
    for (int j = 0; j < 8; j++) take(k + j, r[j]);
  }
  for (; k < k1; k++) take(k, a.chunk_rec[c0 + k]);
  return acc;
}

// "__main__" in a file whose path ends in "py": the per-directory content bit and the lowest such file id
__device__ __forceinline__ void note_main_file(uint32_t f, const uint32_t* file_entry, const uint32_t* entry_dir, const uint32_t* entry_name_off,
                                               const uint8_t* names, uint32_t* dir_content, uint32_t* dir_main) {
  uint32_t e = file_entry[f];
  uint32_t off = entry_name_off[e], nl = entry_name_off[e + 1] - off;
  if (nl >= 2 && names[off + nl - 2] == 'p' && names[off + nl - 1] == 'y') {  // awk '/.py$/' on the path
    // system scope: in a multi-GPU pass these arrays are the ROOT's, reached over NVLink (DevTree::link)
    uint32_t d = entry_dir[e];
    atomicOr_system(&dir_content[d], 1u << M2K_C_MAIN_PY);
    atomicMin_system(&dir_main[d], f);
  }
}

__device__ __forceinline__ void note_dockerfile(const FileArgs& a, uint32_t f) {
  atomicOr_system(&a.dir_content[a.entry_dir[a.file_entry[f]]], 1u << M2K_C_DOCKERFILE);
}

// i: index of the file in the shard.  A file all of whose chunks are here gets its final HAS_MAIN /
// LONG_LINE flags; the part of a file that continues on another GPU becomes a piece summary for
// k_merge_pieces instead.  Nearly every file was already ruled out as a Dockerfile by the scan
// kernel's look at its first word; the others are listed for k_df_head.
__device__ __forceinline__ void finish_file(const FileArgs& a, uint32_t i, uint32_t f, const RunAcc& tot, uint64_t len,
                                            uint32_t has_main, uint32_t not_df) {
  const bool cut_front = i == 0 && a.first_skip != 0;
  const bool cut_back = i + 1 == a.n_shard_files && a.last_end != M2K_NONE;
  if (cut_front || cut_back) {
    PieceSummary ps;
    ps.file = f;
    ps.flags = (has_main ? PIECE_HAS_MAIN : 0u) | (tot.longf ? PIECE_LONG : 0u) | (cut_front ? 0u : PIECE_FIRST) | (not_df ? PIECE_NOT_DF : 0u);
    ps.bytes = tot.bytes; ps.lead = tot.lead; ps.trail = tot.trail;
    a.pieces[i == 0 ? 0 : 1] = ps;
    a.file_flags[f] = 0;
    if (!cut_front && !not_df) { const uint32_t h = atomicAdd(&a.ctr->n_heads, 1u); if (M2K_CHECK(h < a.n_shard_files, 2)) a.head_list[h] = i; }  // this part starts the file: its head can be looked at
    return;
  }
  const uint32_t longf = acc_long(tot.bytes, tot.lead, tot.trail, tot.longf);
  a.file_flags[f] = (uint8_t)((has_main ? M2K_FILE_HAS_MAIN : 0u) | (longf ? M2K_FILE_LONG_LINE : 0u));
  if (has_main) note_main_file(f, a.file_entry, a.entry_dir, a.entry_name_off, a.names, a.dir_content, a.dir_main);
  if (!not_df && !longf && len) { const uint32_t h = atomicAdd(&a.ctr->n_heads, 1u); if (M2K_CHECK(h < a.n_shard_files, 2)) a.head_list[h] = i; }  // a long line fails the parse anyway
}

__global__ void __launch_bounds__(256) k_file_combine(FileArgs a) {
  __shared__ uint32_t s_big[256], s_nbig, s_has[2];
  __shared__ RunAcc s_part[8];
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t i = blockIdx.x * blockDim.x + tid;
  const bool valid = i < a.n_shard_files;
  uint32_t f = 0, c0 = 0, nc = 0;
  uint64_t len = 0, byte0 = 0;
  if (tid == 0) s_nbig = 0;
  __syncthreads();
  if (valid) {
    f = a.file_begin + i;
    len = a.file_len[f];
    c0 = a.chunk_first[i];
    nc = a.chunk_first[i + 1] - c0;
    byte0 = i == 0 ? (uint64_t)a.first_skip * kChunkBytes : 0;
  }
  if (valid && nc <= kCombineWarpChunks) {  // the common case: one thread, a handful of records
    uint32_t has_main = 0, not_df = 0;
    RunAcc tot = fold_records(a, c0, 0, nc, byte0, len, &has_main, &not_df);
    finish_file(a, i, f, tot, len, has_main, not_df);
  } else if (valid) {
    s_big[atomicAdd(&s_nbig, 1u)] = tid;
  }
  __syncthreads();
  // files with many chunks (config D's multi-GiB files: 65 536 records each): the whole block folds them,
  // one at a time -- 256 contiguous shares, merged in order through the warps
  const uint32_t n_big = s_nbig;
  for (uint32_t b = 0; b < n_big; b++) {
    const uint32_t bi = blockIdx.x * blockDim.x + s_big[b], bf = a.file_begin + bi;
    const uint64_t blen = a.file_len[bf], bbyte0 = bi == 0 ? (uint64_t)a.first_skip * kChunkBytes : 0;
    const uint32_t bc0 = a.chunk_first[bi], bnc = a.chunk_first[bi + 1] - bc0;
    const uint32_t per = (bnc + 255) / 256;
    const uint32_t k0 = tid * per < bnc ? tid * per : bnc, k1 = (tid + 1) * per < bnc ? (tid + 1) * per : bnc;
    uint32_t has_main = 0, not_df = 0;
    RunAcc acc = fold_records(a, bc0, k0, k1, bbyte0, blen, &has_main, &not_df);
    if (tid < 2) s_has[tid] = 0;
    RunAcc tot{0, 0, 0, 0};
    for (int l = 0; l < 32; l++) {  // the warp's 32 shares, in order, on lane 0
      RunAcc x;
      x.bytes = __shfl_sync(FULL, acc.bytes, l);
      x.lead = __shfl_sync(FULL, acc.lead, l);
      x.trail = __shfl_sync(FULL, acc.trail, l);
      x.longf = __shfl_sync(FULL, acc.longf, l);
      if (lane == 0) tot = acc_merge(tot, x);
    }
    if (lane == 0) s_part[warp] = tot;
    __syncthreads();
    if (has_main) atomicOr(&s_has[0], 1u);
    if (not_df) atomicOr(&s_has[1], 1u);
    __syncthreads();
    if (tid == 0) {
      RunAcc all{0, 0, 0, 0};
      for (int w = 0; w < 8; w++) all = acc_merge(all, s_part[w]);
      finish_file(a, bi, bf, all, blen, s_has[0], s_has[1]);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// Kernel 4: isDockerFile (dockerfile_dev.cuh) in three launches, none of which ever has a thread walk
// a whole file, and all of which work on dense lists (grids are multiples of the SM count and read
// their sizes from device memory, so the host never waits for a count):
//   k_df_head     one thread per listed file: directives, comments, ARGs, first instruction.
//                 FROM-first files become candidates; a candidate's number and the number of its
//                 first body stripe (kDfStripe = 128 bytes) come from ONE 64-bit counter (high / low half), so the
//                 stripe offsets grow with the candidate number and a stripe finds its candidate
//                 by bisection -- no prefix-sum launch.
//   k_df_states   one thread per stripe, in tiles of 256 consecutive stripes (whatever candidates
//                 they belong to) drawn from a counter: the composed line-state map of each stripe
//                 (does a physical line start a logical line?  one bit carried from the line
//                 before), a segmented block scan, and a decoupled look-back over the tiles before
//                 it for the state a body that began in an earlier tile arrives with.  A tile
//                 publishes its map at once and its end state as soon as it knows it (tagged with
//                 the pass number: the table is never cleared); tickets are handed out in order and
//                 a block holds one only while it runs, so the tile a block waits for is always
//                 held by a running block.  A body of hundreds of MB costs its tiles a few flag
//                 reads each: nothing walks it serially.
//   k_df_validate one thread per stripe: ENV / LABEL / ONBUILD syntax of the logical lines that
//                 start in it; the thread that completes a candidate's last stripe sets its
//                 DOCKERFILE bit and the directory's content bit.
// ---------------------------------------------------------------------------------------------
// Work items of the two parsing kernels below are handed out so that as few of them as possible share a warp:
// threads of one warp that parse different files or stripes diverge, and a diverged warp runs its lanes' paths one
// after the other -- on a small tree (config B: 128 heads, 1551 stripes) that serial run IS the kernel's duration.
// With n items and W warps in the grid, each warp takes k = the power of two >= n / W (at most 32) consecutive
// items per round on its first k lanes; k = 32 is the plain one-item-per-thread order large trees get.
struct Spread { uint32_t k, item0, per_round; bool lane_on; };
__device__ __forceinline__ Spread spread_items(uint32_t n) {
  const uint32_t W = gridDim.x * (blockDim.x >> 5), gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  uint32_t k = 1;
  while (k < 32 && (uint64_t)k * W < n) k <<= 1;
  return Spread{k, gw * k + lane, W * k, lane < k};
}

__global__ void __launch_bounds__(128) k_df_head(FileArgs a) {
  const uint32_t n = a.ctr->n_heads;
  const Spread sp = spread_items(n);
  if (!sp.lane_on) return;
  for (uint32_t j = sp.item0; j < n; j += sp.per_round) {
    const uint32_t i = a.head_list[j], f = a.file_begin + i;
    const uint8_t* fbytes = a.bytes + (a.file_off[f] - a.shard_byte0);
    const uint64_t len = a.file_len[f];
    uint64_t body = 0;
    uint32_t esc = '\\';
    if (i + 1 == a.n_shard_files && a.last_end != M2K_NONE) {
      // only the first part of this file is here: settle what its head can settle; the body of a
      // FROM-first file cannot be validated from a piece (see k_merge_pieces)
      const uint64_t n_here = (uint64_t)a.last_end * kChunkBytes;  // < len
      bool touched_end = false;
      const bool from = df::head(fbytes, n_here, false, &body, &esc, &touched_end);
      a.pieces[i == 0 ? 0 : 1].flags |= touched_end ? PIECE_HEAD_TRUNC : (from ? PIECE_HEAD_FROM : PIECE_HEAD_NOT_FROM);
      continue;
    }
    if (!df::head(fbytes, len, false, &body, &esc)) continue;
    uint32_t flags = a.file_flags[f] | M2K_FILE_FROM_FIRST;
    const uint64_t n_stripes = (len - body + df::kStripe - 1) / df::kStripe;
    if (n_stripes == 0) {  // nothing after the FROM line: nothing left that could fail to parse
      flags |= M2K_FILE_DOCKERFILE;
      note_dockerfile(a, f);
    } else {
      const unsigned long long old = atomicAdd(&a.ctr->cand_stripes, (1ull << 32) | n_stripes);
      const uint32_t c = (uint32_t)(old >> 32);
      if (!M2K_CHECK(c < a.n_shard_files && (uint32_t)old + n_stripes <= a.stripe_cap, 2)) continue;
      a.cand_file[c] = f; a.cand_body[c] = body; a.cand_esc[c] = (uint8_t)esc;
      a.cand_stripe0[c] = (uint32_t)old; a.cand_err[c] = 0; a.cand_done[c] = 0;
    }
    a.file_flags[f] = (uint8_t)flags;
  }
}

__device__ __forceinline__ uint32_t ld_acquire(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// stripe s: its candidate, and the byte range [lo, hi) of the body it covers
struct StripeRef { uint32_t c, k; uint64_t body, lo, hi; };
__device__ __forceinline__ StripeRef df_stripe(const FileArgs& a, uint32_t n_cand, uint32_t s) {
  uint32_t c = 0, h = n_cand;  // the last candidate whose first stripe is <= s
  while (h - c > 1) { uint32_t m = (c + h) >> 1; if (a.cand_stripe0[m] <= s) c = m; else h = m; }
  StripeRef r;
  r.c = c; r.k = s - a.cand_stripe0[c];
  r.body = a.cand_body[c];
  r.lo = r.body + (uint64_t)r.k * df::kStripe;
  const uint64_t n = a.file_len[a.cand_file[c]];
  r.hi = r.lo + df::kStripe < n ? r.lo + df::kStripe : n;
  return r;
}
__device__ __forceinline__ df::File df_file(const FileArgs& a, uint32_t c) {
  const uint32_t f = a.cand_file[c];
  return df::File{a.bytes + (a.file_off[f] - a.shard_byte0), a.file_len[f], a.cand_esc[c]};
}

// segmented scan element: a line-state map and "a body starts at or before this stripe inside the span"
__device__ __forceinline__ uint32_t seg_join(uint32_t first, uint32_t then) {  // bit 2 = segment flag, bits 0-1 = map
  if (then & 4u) return then;
  return (first & 4u) | df::map_compose(first & 3u, then & 3u);
}

// line-state map of every stripe (bits 0-1) and "a body starts here" (bit 2), left in stripe_state for the scan below:
// the parsing half of the state computation, apart from the scan so that its items can be spread over the warps
__global__ void __launch_bounds__(128) k_df_maps(FileArgs a) {
  const unsigned long long cs = a.ctr->cand_stripes;  // final: k_df_head has finished
  const uint32_t n_cand = (uint32_t)(cs >> 32), n_stripes = (uint32_t)cs;
  const Spread sp = spread_items(n_stripes);
  if (!sp.lane_on) return;
  for (uint32_t s = sp.item0; s < n_stripes; s += sp.per_round) {
    const StripeRef r = df_stripe(a, n_cand, s);
    const uint32_t el = df::stripe_map(df_file(a, r.c), r.body, r.lo, r.hi) | (r.k == 0 ? 4u : 0u);
    if (M2K_CHECK(s < a.stripe_cap, 3)) a.stripe_state[s] = (uint8_t)el;
  }
}

__global__ void __launch_bounds__(kDfTileStripes) k_df_states(FileArgs a) {
  __shared__ uint32_t s_ticket, s_in;
  __shared__ uint32_t s_warp[kDfTileStripes / 32];
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t n_stripes = (uint32_t)a.ctr->cand_stripes;  // low half; final: k_df_head has finished
  const uint32_t n_tiles = (n_stripes + kDfTileStripes - 1) / kDfTileStripes;
  const uint32_t tag = a.epoch << 3;
  for (;;) {
    if (tid == 0) s_ticket = atomicAdd(&a.ctr->next_tile, 1u);
    __syncthreads();
    const uint32_t t = s_ticket;
    if (t >= n_tiles) break;
    const uint32_t s = t * kDfTileStripes + tid;
    const bool live = s < n_stripes;
    uint32_t el = 0x2u;  // identity map, no segment start
    if (live && M2K_CHECK(s < a.stripe_cap, 3)) el = a.stripe_state[s];  // k_df_maps left it there
    uint32_t inc = el;  // inclusive segmented scan over the warp (earlier stripe first)
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(FULL, inc, d);
      if ((int)lane >= d) inc = seg_join(o, inc);
    }
    uint32_t exc = __shfl_up_sync(FULL, inc, 1);
    if (lane == 0) exc = 0x2u;
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t pre = 0x2u;
    for (uint32_t w = 0; w < warp; w++) pre = seg_join(pre, s_warp[w]);
    exc = seg_join(pre, exc);  // everything before this stripe in the tile
    if (warp == 0) {
      uint32_t agg = 0x2u;
      for (uint32_t w = 0; w < kDfTileStripes / 32; w++) agg = seg_join(agg, s_warp[w]);
      // the state the tile's first stripe starts in: 0 when a body starts right there, else what the
      // tiles before it leave behind
      const bool starts_body = (__shfl_sync(FULL, el, 0) & 4u) != 0;  // lane 0 of warp 0 is stripe 0 of the tile
      uint32_t sin = 0;
      if (agg & 4u) {
        // a body starts inside this tile: what the tile leaves behind does not depend on what came in
        if (lane == 0) st_release(a.tile_state + t, tag | 4u | df::map_apply(agg & 3u, 0u));
      } else if (lane == 0) {
        st_release(a.tile_state + t, tag | (agg & 3u));  // bit 2 clear: a map, the end state is not known yet
      }
      if (!starts_body && t != 0) {
        uint32_t suffix = 0x2u;  // composed map of the tiles between the window and this one
        for (uint32_t back = 1, done = 0; !done; back += 32) {
          const bool in_range = back + lane <= t;
          uint32_t v = 0;
          if (in_range) while (((v = ld_acquire(a.tile_state + (t - back - lane))) >> 3) != a.epoch) __nanosleep(20);
          const uint32_t known = __ballot_sync(FULL, in_range && (v & 4u));  // nearest tile (lowest lane) whose end state is known
          const uint32_t stop = known ? (uint32_t)(__ffs(known) - 1) : 32u;
          uint32_t window = 0x2u;  // maps of lanes stop-1 .. 0 in file order (farther tile first)
          for (int l = 31; l >= 0; l--) {
            const uint32_t mv = __shfl_sync(FULL, v, l);
            if ((uint32_t)l < stop && back + (uint32_t)l <= t) window = df::map_compose(window, mv & 3u);
          }
          suffix = df::map_compose(window, suffix);
          if (known) { sin = df::map_apply(suffix, __shfl_sync(FULL, v, stop) & 1u); done = 1; }
          else if (back + 32 > t) { sin = df::map_apply(suffix, 0u); done = 1; }  // cannot happen: tile 0 starts a body
        }
      }
      if (lane == 0) {
        if (!(agg & 4u)) st_release(a.tile_state + t, tag | 4u | df::map_apply(agg & 3u, sin));
        s_in = sin;
      }
    }
    __syncthreads();
    // this stripe's own start state: from the start of its body when that lies in the tile, else from the tile's
    if (live && M2K_CHECK(s < a.stripe_cap && t < a.tile_cap, 3)) a.stripe_state[s] = (el & 4u) ? (uint8_t)0 : (uint8_t)df::map_apply(exc & 3u, (exc & 4u) ? 0u : s_in);
  }
}

__global__ void __launch_bounds__(128) k_df_validate(FileArgs a) {
  const unsigned long long cs = a.ctr->cand_stripes;
  const uint32_t n_cand = (uint32_t)(cs >> 32), n_stripes = (uint32_t)cs;
  const uint32_t lane = threadIdx.x & 31;
  const Spread sp = spread_items(n_stripes);
  for (uint32_t base = 0; base < n_stripes; base += sp.per_round) {  // whole warps stay together for the vote below
    const uint32_t s = base + sp.item0;
    const bool live = sp.lane_on && s < n_stripes;
    uint32_t c = M2K_NONE;
    if (live) {
      const StripeRef r = df_stripe(a, n_cand, s);
      c = r.c;
      (void)M2K_CHECK(r.lo >= r.body && r.lo < r.hi && r.hi <= a.file_len[a.cand_file[c]], 4);
      if (!df::stripe_validate(df_file(a, c), r.body, r.lo, r.hi, a.stripe_state[s])) {
        atomicExch(&a.cand_err[c], 1u);
        __threadfence();  // visible before this warp's leader counts the stripe as finished (the vote below orders the two)
      }
    }
    // count finished stripes per candidate, one atomic per (warp, candidate); whoever completes a
    // candidate publishes its verdict
    const uint32_t peers = __match_any_sync(FULL, c);
    if (live && (uint32_t)(__ffs(peers) - 1) == lane) {
      const uint32_t mine = (uint32_t)__popc(peers);
      const uint32_t total = (c + 1 < n_cand ? a.cand_stripe0[c + 1] : n_stripes) - a.cand_stripe0[c];
      __threadfence();
      if (atomicAdd(&a.cand_done[c], mine) + mine == total) {
        __threadfence();
        if (atomicAdd(&a.cand_err[c], 0u) == 0u) {
          const uint32_t f = a.cand_file[c];
          a.file_flags[f] |= M2K_FILE_DOCKERFILE;
          note_dockerfile(a, f);
        }
      }
    }
  }
}

static FileArgs file_args(DevTree* t, const ChunkRec* chunk_rec) {
  // a rank that is attached to a root reduces its per-directory contributions straight into the root's arrays
  const bool remote = t->link.attached && !t->link.is_root;
  return FileArgs{t->file_end - t->file_begin, t->file_begin, t->first_skip, t->last_end, t->chunk_first, chunk_rec,
                  t->file_off, t->file_len, t->bytes, t->shard_byte0, t->file_entry, t->entry_dir, t->entry_name_off, t->names,
                  t->file_flags, remote ? t->link.dir_content : t->dir_content, remote ? t->link.dir_main : t->dir_main, t->pieces, t->df_counters, t->head_list, t->cand_file, t->cand_body,
                  t->cand_esc, t->cand_stripe0, t->cand_err, t->cand_done, t->stripe_state, t->tile_state, t->epoch, t->stripe_cap, t->tile_cap};
}

// per-file combine, then isDockerFile: heads, stripe states, validation
void launch_file_phase(Ctx* c, DevTree* t, const ChunkRec* chunk_rec) {
  uint32_t nf = t->file_end - t->file_begin;
  cudaMemsetAsync(t->pieces, 0xFF, 2 * sizeof(PieceSummary), c->stream);  // file == M2K_NONE: unused slot
  if (!nf) return;
  FileArgs a = file_args(t, chunk_rec);
  k_file_combine<<<(nf + 255) / 256, 256, 0, c->stream>>>(a);
  c->launches++;
}
void launch_dockerfile(Ctx* c, DevTree* t) {
  if (t->file_end == t->file_begin) return;
  FileArgs a = file_args(t, nullptr);
  k_df_head<<<c->sm_count * 8, 128, 0, c->stream>>>(a);
  k_df_maps<<<c->sm_count * 8, 128, 0, c->stream>>>(a);
  k_df_states<<<c->sm_count * 4, kDfTileStripes, 0, c->stream>>>(a);
  k_df_validate<<<c->sm_count * 8, 128, 0, c->stream>>>(a);
  c->launches += 4;
}

// ---------------------------------------------------------------------------------------------
// Kernel 5: what the byte scan adds to the per-directory results.  The content results are sparse
// (directories holding a python main file or a Dockerfile): each such directory pushes its bits and
// its lowest main-file id up its ancestor chain, and stops where nothing new arrives (the thread
// that brought the same or better continues from there).
// ---------------------------------------------------------------------------------------------
__global__ void k_content_up(uint32_t n_dirs, const uint32_t* __restrict__ dir_parent, const uint32_t* __restrict__ dir_content,
                             const uint32_t* __restrict__ dir_main, uint32_t* __restrict__ dir_subtree, uint32_t* __restrict__ dir_main_sub) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n_dirs) return;
  uint32_t bits = dir_content[d], mf = dir_main[d];
  for (uint32_t x = d; bits || mf != M2K_NONE; x = dir_parent[x]) {
    if (bits && (atomicOr(&dir_subtree[x], bits) & bits) == bits) bits = 0;
    if (mf != M2K_NONE && atomicMin(&dir_main_sub[x], mf) <= mf) mf = M2K_NONE;
    if (x == 0) break;
  }
}
void launch_content_up(Ctx* c, DevTree* t, cudaStream_t st) {
  if (!t->n_dirs) return;
  k_content_up<<<(t->n_dirs + 255) / 256, 256, 0, st>>>(t->n_dirs, t->dir_parent, t->dir_content, t->dir_main, t->dir_subtree, t->dir_main_sub);
  c->launches++;
}

// ---------------------------------------------------------------------------------------------
// Kernel 6: the 18 built-in predicates of a directory (SURVEY.md App. A truth table), evaluated inside
// the walk kernel below.  D = features of the directory's own entries, S = OR over the subtree.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t dir_options(uint32_t D, uint32_t S, uint32_t W, uint32_t sh_flavor) {
  auto F = [&](int b) { return (D >> b) & 1u; };
  auto R = [&](int b) { return (S >> b) & 1u; };
  uint32_t py = (sh_flavor == M2K_SH_BASH) ? (F(M2K_F_REQUIREMENTS) | F(M2K_F_SETUP_PY) | F(M2K_F_ENVIRONMENT_YML) | F(M2K_F_PIPFILE)) : 0u;
  uint32_t m = 0;
  m |= F(M2K_F_PIPFILE) << M2K_DF_DJANGO;
  m |= R(M2K_R_GO) << M2K_DF_GOLANG;
  m |= (W << M2K_DF_WAR_JBOSS) | (W << M2K_DF_WAR_LIBERTY) | (W << M2K_DF_WAR_TOMCAT);
  m |= F(M2K_F_BUILD_XML) << M2K_DF_JAVAANT;
  m |= F(M2K_F_BUILD_GRADLE) << M2K_DF_JAVAGRADLE;
  m |= F(M2K_F_POM_XML) << M2K_DF_JAVAMAVEN;
  m |= F(M2K_F_PACKAGE_JSON) << M2K_DF_NODEJS;
  m |= R(M2K_R_PHP) << M2K_DF_PHP;
  m |= py << M2K_DF_PYTHON;
  m |= F(M2K_F_GEMFILE) << M2K_DF_RUBY;
  m |= (F(M2K_F_GO_MOD) | R(M2K_R_GO)) << M2K_S2I_GOLANG;
  m |= ((1u ^ F(M2K_F_BUILD_GRADLE)) & (1u ^ F(M2K_F_BUILD_XML)) & (F(M2K_F_POM_XML) | R(M2K_R_JAVA))) << M2K_S2I_JAVA;
  m |= F(M2K_F_PACKAGE_JSON) << M2K_S2I_NODEJS;
  m |= R(M2K_R_PHP) << M2K_S2I_PHP;
  m |= py << M2K_S2I_PYTHON;
  m |= F(M2K_F_GEMFILE) << M2K_S2I_RUBY;
  return m;
}

// ---------------------------------------------------------------------------------------------
// Kernel 7: the walk (any2kube.go:58-98) in two launches.  k_walk_local: per directory, its option
// bits (kernel 6), the case-folded hash of its path, a probe of the ignore table (an exact hash join in place of
// O(#dirs x #entries) EqualFold scans, utils.go:308-315; a hash hit is confirmed rune by rune against
// the folded entry, so 64-bit collisions cannot change a decision) and the walk's local rule.
// k_walk_reach: a directory is visited iff the walk descends below all of its ancestors.
// ---------------------------------------------------------------------------------------------
struct WalkArgs {
  const uint32_t* dir_parent; const uint32_t* dir_entry; const uint8_t* dir_flags;
  const uint32_t* entry_name_off; const uint8_t* names;
  const uint32_t* fold_from; const uint32_t* fold_to;
  const IgnoreSlot* table; uint32_t table_mask;
  const uint32_t* runes;          // folded runes of all table entries
  const uint32_t* prefix_runes;   // folded runes of the path prefix shared by depth-1 directories
  uint32_t prefix_len;
  uint64_t depth1_state; uint32_t root_ignore;
  const uint32_t* dir_direct; const uint32_t* dir_subtree; const uint32_t* dir_first_war; const uint8_t* entry_type;
  uint32_t sh_flavor;
  uint32_t* dir_opts;
  uint8_t* dir_ignore; uint8_t* dir_state; uint8_t* dir_descend;
};

// does the folded path of directory d equal entry runes E[0..L)?
__device__ bool path_equals_entry(const WalkArgs& a, uint32_t d, const uint32_t* E, uint32_t L) {
  uint32_t pos = L;
  for (uint32_t cur = d;;) {
    uint32_t e = a.dir_entry[cur];
    const uint8_t* nm = a.names + a.entry_name_off[e];
    uint32_t len = a.entry_name_off[e + 1] - a.entry_name_off[e];
    uint32_t nr = 0;
    for (uint32_t i = 0; i < len;) { uint32_t w; decode_rune(nm + i, len - i, &w); i += w; nr++; }
    if (pos < nr) return false;
    uint32_t k = pos - nr;
    for (uint32_t i = 0; i < len;) {
      uint32_t w, r = decode_rune(nm + i, len - i, &w);
      if (E[k++] != fold_rune(r, a.fold_from, a.fold_to, M2K_CASEFOLD_N)) return false;
      i += w;
    }
    pos -= nr;
    uint32_t par = a.dir_parent[cur];
    if (par == 0) break;
    if (pos == 0 || E[pos - 1] != '/') return false;
    pos--;
    cur = par;
  }
  if (pos != a.prefix_len) return false;
  for (uint32_t i = 0; i < pos; i++) if (E[i] != a.prefix_runes[i]) return false;
  return true;
}

constexpr uint32_t kChainMax = 64;

// folded hash of "/name" appended to h (or of "name" right after the shared prefix)
__device__ __forceinline__ uint64_t hash_dir_name(const WalkArgs& a, uint64_t h, uint32_t d, bool first) {
  if (!first) h = hash_step(h, '/');
  uint32_t e = a.dir_entry[d];
  uint32_t off = a.entry_name_off[e];
  return hash_fold_bytes(h, a.names + off, a.entry_name_off[e + 1] - off, a.fold_from, a.fold_to, M2K_CASEFOLD_N);
}

// Step 1, one thread per directory, no dependency between directories: the 18 predicates, the
// case-folded hash of the directory's own path (hashing down its ancestor chain), the ignore-table
// probe with exact confirmation, and the walk's LOCAL rule -- what any2kube.go:58-98 does with this
// directory if it is reached, and whether the walk would descend below it.
__global__ void __launch_bounds__(128) k_walk_local(uint32_t n_dirs, const uint32_t* __restrict__ dir_depth, WalkArgs a) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n_dirs) return;
  const uint32_t fw = a.dir_first_war[d];
  const uint32_t W = (fw != M2K_NONE && (a.entry_type[fw] >> 4) != M2K_T_ERR) ? 1u : 0u;  // [ -e first match ]
  const uint32_t opts = dir_options(a.dir_direct[d], a.dir_subtree[d], W, a.sh_flavor);
  a.dir_opts[d] = opts;
  uint32_t ign = 0;
  if (d == 0) {
    ign = a.root_ignore;  // the root path is compared on the host exactly as given
  } else if (a.table_mask != M2K_NONE) {
    const uint32_t depth = dir_depth[d];
    uint64_t h = a.depth1_state;
    if (depth <= kChainMax) {
      uint32_t chain[kChainMax];
      uint32_t x = d;
      for (uint32_t i = depth; i-- > 0;) { chain[i] = x; x = a.dir_parent[x]; }
      for (uint32_t i = 0; i < depth; i++) h = hash_dir_name(a, h, chain[i], i == 0);
    } else {
      // very deep paths: find the ancestor of each level by walking up again (quadratic, but only here)
      for (uint32_t lvl = 1; lvl <= depth; lvl++) {
        uint32_t x = d;
        for (uint32_t up = depth - lvl; up > 0; up--) x = a.dir_parent[x];
        h = hash_dir_name(a, h, x, lvl == 1);
      }
    }
    uint64_t key = hash_finish(h);
    const IgnoreSlot* region = a.table + (size_t)(key >> (64 - kIgnoreRegionBits)) * (a.table_mask + 1);
    for (uint32_t s = (uint32_t)key & a.table_mask;; s = (s + 1) & a.table_mask) {
      IgnoreSlot sl = region[s];
      if (sl.hash == 0) break;
      if (sl.hash == key && path_equals_entry(a, d, a.runes + sl.off, slot_len(sl.len_flags))) { ign = slot_flags(sl.len_flags); break; }
    }
  }
  a.dir_ignore[d] = (uint8_t)ign;
  uint32_t state, descend = 0;
  if (a.dir_flags[d] & 1) state = M2K_DIR_WALK_ERROR;
  else if (ign & 4) state = M2K_DIR_PRECONTAINERIZED;
  else if (ign & 1) { if (ign & 2) state = M2K_DIR_IGNORED_PRUNE; else { state = M2K_DIR_IGNORED_DESCEND; descend = 1; } }
  else if (opts == 0) { if (ign & 2) state = M2K_DIR_NOMATCH_PRUNE; else { state = M2K_DIR_NOMATCH_DESCEND; descend = 1; } }
  else state = M2K_DIR_SERVICE;
  a.dir_state[d] = (uint8_t)state;
  a.dir_descend[d] = (uint8_t)descend;
}

// Step 2: a directory is reached iff the walk descends below every one of its ancestors.
__global__ void __launch_bounds__(128) k_walk_reach(uint32_t n_dirs, WalkArgs a) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n_dirs || d == 0) return;
  for (uint32_t x = a.dir_parent[d];; x = a.dir_parent[x]) {
    if (!a.dir_descend[x]) { a.dir_state[d] = M2K_DIR_NOT_VISITED; return; }
    if (x == 0) return;
  }
}

void launch_walk(Ctx* c, DevTree* t, uint32_t sh_flavor, const IgnoreSlot* table, uint32_t table_mask, const uint32_t* runes,
                 const uint32_t* prefix_runes, uint32_t prefix_len, uint64_t depth1_state, uint8_t root_ignore, cudaStream_t st) {
  if (!t->n_dirs) return;
  WalkArgs a{t->dir_parent, t->dir_entry, t->dir_flags, t->entry_name_off, t->names, c->d_fold_from, c->d_fold_to,
             table, table_mask, runes, prefix_runes, prefix_len, depth1_state, root_ignore,
             t->dir_direct, t->dir_subtree, t->dir_first_war, t->entry_type, sh_flavor, t->dir_opts,
             t->dir_ignore, t->dir_state, t->dir_descend};
  uint32_t blocks = (t->n_dirs + 127) / 128;
  k_walk_local<<<blocks, 128, 0, st>>>(t->n_dirs, t->dir_depth, a);
  k_walk_reach<<<blocks, 128, 0, st>>>(t->n_dirs, a);
  c->launches += 2;
}

const void* k_merge_partials_fn();
const void* k_merge_pieces_fn();
const void* k_settle_file_fn();
const void* k_synth_fill_fn();

// ---------------------------------------------------------------------------------------------
// multi-GPU over peer memory (PeerLink): pass ordering flags in the root's mailbox, and what a rank
// sends besides its in-kernel reductions -- its own range of the per-file flags and its two piece
// summaries, written into the root's memory over NVLink.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__global__ void k_signal(uint32_t* flag, uint32_t value) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  __threadfence_system();  // everything this stream did before (kernel boundaries) is visible before the flag
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}
// one thread per flag waits until it reaches `value` (pass numbers only grow).  A peer that never
// arrives must not hang the GPU: after 30 s the wait gives up and raises `abort`, which the host
// reports as an error of the pass.
__global__ void k_wait_ge(const uint32_t* flags, uint32_t n_flags, uint32_t value, uint32_t* abort) {
  if (threadIdx.x >= n_flags || blockIdx.x != 0) return;
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  while ((int32_t)(ld_acquire_sys(flags + threadIdx.x) - value) < 0) {
    __nanosleep(200);
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > 30000000000ull) { atomicExch_system(abort, 1u); break; }
  }
  __threadfence_system();
}
__global__ void __launch_bounds__(256) k_contribute(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint32_t lo, uint32_t hi,
                                                    const PieceSummary* __restrict__ my_pieces, PieceSummary* __restrict__ root_pieces) {
  // bytes [lo, hi) of the per-file flags: both arrays are 256-byte aligned, so index alignment is address alignment
  const uint32_t a_lo = (lo + 15) & ~15u, a_hi = hi & ~15u;
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  if (a_lo < a_hi) {
    for (uint32_t i = a_lo / 16 + tid; i < a_hi / 16; i += nth) reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
    for (uint32_t i = lo + tid; i < a_lo; i += nth) dst[i] = src[i];
    for (uint32_t i = a_hi + tid; i < hi; i += nth) dst[i] = src[i];
  } else {
    for (uint32_t i = lo + tid; i < hi; i += nth) dst[i] = src[i];
  }
  if (tid < 2) root_pieces[tid] = my_pieces[tid];
}

// Kernels are loaded lazily, on first launch, and loading can wait for the device to go idle: a first
// launch made while a k_wait_ge spins for the very work that launch belongs to would never return.
// Every kernel a pass can launch is therefore loaded when an exchange is set up.
int preload_kernels() {
  const void* fns[] = {(const void*)k_fill_u32, (const void*)k_name_features, (const void*)k_scan_bytes, (const void*)k_scan_bytes_bulk,
                       (const void*)k_file_combine, (const void*)k_df_head, (const void*)k_df_maps, (const void*)k_df_states, (const void*)k_df_validate,
                       (const void*)k_content_up, (const void*)k_walk_local, (const void*)k_walk_reach, (const void*)k_signal,
                       (const void*)k_wait_ge, (const void*)k_contribute, (const void*)k_merge_partials_fn(), (const void*)k_merge_pieces_fn(),
                       (const void*)k_settle_file_fn(), (const void*)k_synth_fill_fn()};
  for (const void* f : fns) {
    cudaFuncAttributes attr;
    cudaError_t e = cudaFuncGetAttributes(&attr, f);
    if (e != cudaSuccess) return (int)e;
  }
  return 0;
}

void launch_signal(Ctx* c, uint32_t* flag, uint32_t value, cudaStream_t st) {
  k_signal<<<1, 32, 0, st>>>(flag, value);
  c->launches++;
}
void launch_wait_ge(Ctx* c, const uint32_t* flags, uint32_t n_flags, uint32_t value, uint32_t* abort, cudaStream_t st) {
  if (!n_flags) return;
  k_wait_ge<<<1, 64, 0, st>>>(flags, n_flags, value, abort);
  c->launches++;
}
// a non-root rank, after its per-file kernels: its flag range and piece summaries into the root's memory
void launch_contribute(Ctx* c, DevTree* t, cudaStream_t st) {
  const uint32_t n = t->file_end - t->file_begin;
  uint32_t blocks = (n / 16 + 255) / 256 + 1;
  if (blocks > (uint32_t)c->sm_count * 4) blocks = (uint32_t)c->sm_count * 4;
  k_contribute<<<blocks, 256, 0, st>>>(t->file_flags, t->link.file_flags, t->file_begin, t->file_end, t->pieces,
                                        t->link.pieces + 2 * t->link.rank);
  c->launches++;
}

// ---------------------------------------------------------------------------------------------
// multi-GPU through a collective library (the caller gathers): fold the gathered per-directory
// partials (rank-major) into this rank's arrays
// ---------------------------------------------------------------------------------------------
__global__ void k_merge_partials(uint32_t n_dirs, uint32_t n_ranks, const uint32_t* __restrict__ all_content,
                                 const uint32_t* __restrict__ all_main, uint32_t* __restrict__ dir_content, uint32_t* __restrict__ dir_main) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n_dirs) return;
  uint32_t o = 0, m = M2K_NONE;
  for (uint32_t r = 0; r < n_ranks; r++) {
    o |= all_content[(size_t)r * n_dirs + d];
    uint32_t v = all_main[(size_t)r * n_dirs + d];
    m = v < m ? v : m;
  }
  dir_content[d] = o;
  dir_main[d] = m;
}

void launch_merge_partials(Ctx* c, DevTree* t, const uint32_t* all_content, const uint32_t* all_main, uint32_t n_ranks) {
  if (!t->n_dirs) return;
  k_merge_partials<<<(t->n_dirs + 255) / 256, 256, 0, c->stream>>>(t->n_dirs, n_ranks, all_content, all_main, t->dir_content, t->dir_main);
  c->launches++;
}

// Pieces of files that were scanned on several GPUs, in rank order (slot 0 then slot 1 of each rank;
// unused slots have file == M2K_NONE): consecutive pieces of one file are stitched with the same
// monoid the combine kernel uses.  One thread: there are at most two pieces per rank.
// isDockerFile for such a file: the first word or the head (parsed on the GPU that holds the start of
// the file) or a long line usually settles it; a FROM-first file without a long line would need its
// body validated across GPUs, which no piece can do alone -- it is marked M2K_FILE_UNDECIDED and the
// caller settles it with a pass over that one file (m2k_upload of [f, f+1)).
struct MergePiecesArgs {
  const uint64_t* file_len; const uint32_t* file_entry; const uint32_t* entry_dir; const uint32_t* entry_name_off; const uint8_t* names;
  uint8_t* file_flags; uint32_t* dir_content; uint32_t* dir_main;
};
__global__ void k_merge_pieces(const PieceSummary* __restrict__ all, uint32_t n, MergePiecesArgs a) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  uint32_t k = 0;
  while (k < n) {
    if (all[k].file == M2K_NONE) { k++; continue; }
    const uint32_t f = all[k].file;
    RunAcc tot{0, 0, 0, 0};
    uint32_t flags = 0;
    bool first_seen = false;
    for (; k < n; k++) {
      const PieceSummary ps = all[k];
      if (ps.file == M2K_NONE) continue;
      if (ps.file != f) break;
      if (tot.bytes == 0) first_seen = (ps.flags & PIECE_FIRST) != 0;
      RunAcc b{ps.bytes, ps.lead, ps.trail, (ps.flags & PIECE_LONG) ? 1u : 0u};
      tot = acc_merge(tot, b);
      flags |= ps.flags;
    }
    uint32_t out;
    if (!first_seen || tot.bytes != a.file_len[f]) {
      out = M2K_FILE_UNDECIDED;  // pieces missing (a caller gathered fewer ranks than it sharded over)
    } else {
      const uint32_t longf = acc_long(tot.bytes, tot.lead, tot.trail, tot.longf);
      out = ((flags & PIECE_HAS_MAIN) ? M2K_FILE_HAS_MAIN : 0u) | (longf ? M2K_FILE_LONG_LINE : 0u);
      const bool settled_no = (flags & (PIECE_NOT_DF | PIECE_HEAD_NOT_FROM)) || longf;
      if (!settled_no) out |= M2K_FILE_UNDECIDED;
      if (flags & PIECE_HAS_MAIN) note_main_file(f, a.file_entry, a.entry_dir, a.entry_name_off, a.names, a.dir_content, a.dir_main);
    }
    a.file_flags[f] = (uint8_t)out;
  }
}

__global__ void k_settle_file(uint32_t f, uint32_t flags, const uint32_t* __restrict__ file_entry, const uint32_t* __restrict__ entry_dir,
                              uint8_t* __restrict__ file_flags, uint32_t* __restrict__ dir_content) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  file_flags[f] = (uint8_t)flags;
  if (flags & M2K_FILE_DOCKERFILE) atomicOr(&dir_content[entry_dir[file_entry[f]]], 1u << M2K_C_DOCKERFILE);
}
void launch_settle_file(Ctx* c, DevTree* t, uint32_t f, uint32_t flags) {
  k_settle_file<<<1, 32, 0, c->stream>>>(f, flags, t->file_entry, t->entry_dir, t->file_flags, t->dir_content);
  c->launches++;
}

void launch_merge_pieces(Ctx* c, DevTree* t, const PieceSummary* all_pieces, uint32_t n_pieces) {
  if (!n_pieces) return;
  MergePiecesArgs a{t->file_len, t->file_entry, t->entry_dir, t->entry_name_off, t->names, t->file_flags, t->dir_content, t->dir_main};
  k_merge_pieces<<<1, 32, 0, c->stream>>>(all_pieces, n_pieces, a);
  c->launches++;
}

// ---------------------------------------------------------------------------------------------
// synthetic content, generated in place in HBM (same function as the host fill, synth.h)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_synth_fill(
    uint32_t n_chunks, uint32_t n_shard_files, const uint32_t* __restrict__ chunk_first, const uint64_t* __restrict__ file_off,
    const uint64_t* __restrict__ file_len, uint32_t file_begin, const uint8_t* __restrict__ synth_class,
    const uint64_t* __restrict__ synth_inject, uint64_t seed, uint8_t* __restrict__ bytes, uint64_t shard_byte0,
    uint32_t first_skip, uint32_t last_end) {
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t n_warps = gridDim.x * (blockDim.x / 32);
  for (uint32_t item = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5); item < n_chunks; item += n_warps) {
    uint32_t lo = 0, hi = n_shard_files;
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (chunk_first[mid] <= item) lo = mid; else hi = mid; }
    uint32_t f = file_begin + lo;
    uint64_t flen = file_len[f];
    uint64_t ublock0 = (uint64_t)(item - chunk_first[lo] + (lo == 0 ? first_skip : 0u)) * kUnitsPerChunk;
    uint64_t total_units = (flen + 15) >> 4;
    uint32_t n_units = (uint32_t)(total_units - ublock0 < kUnitsPerChunk ? total_units - ublock0 : kUnitsPerChunk);
    uint64_t key = synth_file_key(seed, f);
    uint8_t cls = synth_class[f];
    uint64_t inj = synth_inject[f];
    uint4* dst = reinterpret_cast<uint4*>(bytes + (file_off[f] - shard_byte0)) + ublock0;
    for (uint32_t u = lane; u < n_units; u += 32) {
      union { uint8_t b[16]; uint4 v; } blk;
      synth_block16(key, cls, flen, inj, ublock0 + u, blk.b);
      dst[u] = blk.v;
    }
    // a shard that is cut through a file also holds the 16 bytes on the far side of each cut
    if (lane == 0 && item == 0 && first_skip != 0) {
      union { uint8_t b[16]; uint4 v; } blk;
      synth_block16(key, cls, flen, inj, ublock0 - 1, blk.b);
      *(reinterpret_cast<uint4*>(bytes + (file_off[f] - shard_byte0)) + (ublock0 - 1)) = blk.v;
    }
    if (lane == 0 && item + 1 == n_chunks && last_end != M2K_NONE) {
      union { uint8_t b[16]; uint4 v; } blk;
      synth_block16(key, cls, flen, inj, ublock0 + n_units, blk.b);
      dst[n_units] = blk.v;
    }
  }
}

void launch_synth_fill(Ctx* c, DevTree* t, uint64_t seed) {
  if (!t->n_chunks) return;
  uint32_t want = (t->n_chunks + 7) / 8;
  uint32_t cap = (uint32_t)c->sm_count * 8 * 4;
  k_synth_fill<<<want < cap ? want : cap, 256, 0, c->stream>>>(t->n_chunks, t->file_end - t->file_begin, t->chunk_first, t->file_off,
                                                                t->file_len, t->file_begin, t->synth_class, t->synth_inject, seed,
                                                                t->bytes, t->shard_byte0, t->first_skip, t->last_end);
  c->launches++;
}

const void* k_merge_partials_fn() { return (const void*)k_merge_partials; }
const void* k_merge_pieces_fn() { return (const void*)k_merge_pieces; }
const void* k_settle_file_fn() { return (const void*)k_settle_file; }
const void* k_synth_fill_fn() { return (const void*)k_synth_fill; }

}  // namespace m2k
