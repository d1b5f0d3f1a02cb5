// Internal declarations shared by the host code and the kernels of libm2kdetect.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/m2kdetect.h"

namespace m2k {

constexpr uint32_t kChunkBytes = 65536;  // bytes of one scan work item (one warp)
constexpr uint32_t kUnitsPerChunk = kChunkBytes / 16;
#ifndef M2K_DF_STRIPE
#define M2K_DF_STRIPE 128
#endif
constexpr uint32_t kDfStripe = M2K_DF_STRIPE;  // bytes of one stripe of the parallel Dockerfile validation
constexpr uint32_t kDfTileStripes = 256;  // stripes per tile (one block of k_df_states)
constexpr uint32_t kLongLine = 65536;    // bufio.MaxScanTokenSize: a line of >= this many bytes fails
// per-chunk record written by the byte-scan kernel: scratch memory, 8 bytes per 64 KiB chunk, exact
// to the byte so that records of one file can be stitched anywhere (another kernel, another GPU)
// without looking at the file again
//   bits  0..16  lead : bytes before the first newline of the chunk (== bytes in the chunk when it has none)
//   bits 17..33  trail: bytes after the last newline of the chunk  (== bytes in the chunk when it has none)
//   bit  34      the chunk contains the start of a "__main__"
//   bit  35      (first chunk of a file) the first word already rules out a Dockerfile
typedef uint64_t ChunkRec;
__host__ __device__ inline ChunkRec rec_pack(uint32_t lead, uint32_t trail, uint32_t has_main, uint32_t not_df) {
  return (ChunkRec)lead | ((ChunkRec)trail << 17) | ((ChunkRec)has_main << 34) | ((ChunkRec)not_df << 35);
}
__host__ __device__ inline uint32_t rec_lead(ChunkRec r) { return (uint32_t)(r & 0x1FFFF); }
__host__ __device__ inline uint32_t rec_trail(ChunkRec r) { return (uint32_t)((r >> 17) & 0x1FFFF); }
__host__ __device__ inline uint32_t rec_main(ChunkRec r) { return (uint32_t)((r >> 34) & 1); }
__host__ __device__ inline uint32_t rec_not_df(ChunkRec r) { return (uint32_t)((r >> 35) & 1); }

// A shard is a range of 64 KiB chunks and may begin and end inside a (large) file.  The at most two
// files it holds only part of are folded into piece summaries instead of final flags; the pieces of
// a file are stitched in rank order after the exchange (k_merge_pieces).  32 bytes, exchanged as is.
struct PieceSummary {
  uint32_t file;         // file id; M2K_NONE = unused slot
  uint32_t flags;        // PIECE_*
  uint64_t bytes;        // content bytes of the file covered by this piece
  uint64_t lead, trail;  // bytes before the first / after the last newline of the piece (== bytes when it has none)
};
enum : uint32_t {
  PIECE_HAS_MAIN = 1, PIECE_LONG = 2,
  PIECE_FIRST = 4,           // the piece starts at byte 0 of the file
  PIECE_NOT_DF = 8,          // the first word rules out a Dockerfile
  PIECE_HEAD_NOT_FROM = 16,  // the head was parsed within this piece: the first instruction is not FROM / parse error
  PIECE_HEAD_FROM = 32,      // ... it is a FROM line: the body must be validated over the whole file
  PIECE_HEAD_TRUNC = 64      // the head runs past this piece
};

// counters of the isDockerFile phase: candidates in the high half and body tiles in the low half of
// ONE word (so both numberings grow together), and the next tile ticket
struct DfCounters { unsigned long long cand_stripes; uint32_t next_tile; uint32_t scan_ticket; uint32_t n_heads; uint32_t _pad; };  // scan_ticket: see ScanArgs::ticket

// name signature table staged into shared memory by the name-feature kernel
struct NameSig {
  uint64_t lo, hi;       // the last <=16 bytes of the pattern, right aligned (byte 0 of `lo` = last character)
  uint64_t mask_lo, mask_hi;
  uint32_t len;          // pattern length
  uint32_t kind;         // 0 exact name, 1 suffix
  uint32_t feature;      // bit to set (M2K_F_* / M2K_R_* / M2K_N_*)
  uint32_t need;         // 0 any type; 1 followed type must be REG; 2 lstat type must not be DIR; 3 non-dot name (war glob)
};
constexpr int kMaxNameSigs = 32;

// the ignore table is 16 independent linear-probing regions selected by the top bits of the key
constexpr uint32_t kIgnoreRegionBits = 4, kIgnoreRegions = 1u << kIgnoreRegionBits;
struct IgnoreSlot { uint64_t hash; uint32_t off; uint32_t len_flags; };  // folded runes [off, off+len); flags in the top 3 bits
__host__ __device__ inline uint32_t slot_pack(uint32_t len, uint32_t flags) { return len | (flags << 29); }
__host__ __device__ inline uint32_t slot_len(uint32_t lf) { return lf & 0x1FFFFFFFu; }
__host__ __device__ inline uint32_t slot_flags(uint32_t lf) { return lf >> 29; }

// ---- multi-GPU exchange over peer memory (NVLink / NVSwitch) ----
// The rank that assembles the result (the root) owns a mailbox next to its per-directory arrays; the
// other ranks map that memory (peer access inside one process, CUDA IPC between processes) and their
// per-file kernels OR / MIN their sparse per-directory contributions straight into the root's arrays
// -- the reduction happens in the root's L2, as the contributions are produced, and no collective
// library is in the data path.  Flags in the mailbox order the passes: `ready` = the root has
// cleared its arrays for pass p; done[r] = rank r's contributions for pass p have landed.
constexpr uint32_t kMaxRanks = 62;
struct Mailbox { uint32_t ready; uint32_t abort; uint32_t done[kMaxRanks]; };  // 256 bytes
struct PeerLink {
  bool attached = false, is_root = false;
  uint32_t rank = 0, n_ranks = 1;
  uint32_t pass = 0;                 // exchange pass number (every rank counts its m2k_scan calls)
  Mailbox* mailbox = nullptr;        // the root's, as seen from this rank's device
  uint32_t *dir_content = nullptr, *dir_main = nullptr;  // the root's arrays
  uint8_t* file_flags = nullptr;
  PieceSummary* pieces = nullptr;    // [2 * n_ranks] on the root, rank order
  void* ipc_base = nullptr;          // mapping opened with cudaIpcOpenMemHandle (other process), to close
};

// a tree (metadata replicated) plus the bytes of file ids [file_begin, file_end) resident in HBM
struct DevTree {
  const m2k_tree* host = nullptr;
  uint32_t n_dirs = 0, n_entries = 0, n_files = 0, max_depth = 0;
  uint32_t file_begin = 0, file_end = 0;  // shard: the files it touches
  uint32_t first_skip = 0;                // chunks of file `file_begin` that precede the shard (on another GPU)
  uint32_t last_end = M2K_NONE;           // end chunk (within the file) of file `file_end - 1` when the shard stops inside it
  PieceSummary* pieces = nullptr;         // [2] device
  uint64_t shard_bytes = 0;               // padded bytes resident
  uint64_t shard_content = 0;             // sum of file_len over the shard
  uint64_t shard_byte0 = 0;               // file_off[file_begin]
  uint32_t n_chunks = 0;
  // metadata
  uint32_t *dir_parent = nullptr, *dir_entry = nullptr, *dir_depth = nullptr;
  uint8_t* dir_flags = nullptr;
  uint32_t *entry_dir = nullptr, *entry_name_off = nullptr, *entry_child = nullptr;
  uint8_t *entry_type = nullptr, *names = nullptr;
  uint32_t* file_entry = nullptr;
  uint64_t *file_off = nullptr, *file_len = nullptr;  // [n_files] (all files; offsets are tree-global)
  uint32_t* chunk_first = nullptr;        // [shard files + 1] prefix of ceil(len / kChunkBytes) over the shard
  uint32_t* chunk_file = nullptr;         // [n_chunks] its inverse: the shard-local file index of every chunk
  uint8_t* bytes = nullptr;               // shard bytes; byte 0 is tree offset shard_byte0
  struct Ctx* arena_ctx = nullptr;        // the context whose arena `bytes` / `stripe_state` came from (bit 0 / 1 of from_arena)
  uint32_t from_arena = 0;
  // synthetic content description (device copies), only for m2k_synth_fill_device
  uint8_t* synth_class = nullptr;
  uint64_t* synth_inject = nullptr;
  // work buffers
  ChunkRec* chunk_rec = nullptr;          // [n_chunks] scratch
  // isDockerFile scratch: the files whose head must be parsed and the FROM-first candidates among
  // them (worst case: every file), one byte per stripe (kDfStripe bytes) of candidate body (worst case: every
  // byte of the shard), one word per tile of 256 stripes (tagged with `epoch`, so never cleared)
  DfCounters* df_counters = nullptr;      // in the slab's zero region
  uint32_t *head_list = nullptr, *cand_file = nullptr, *cand_stripe0 = nullptr, *cand_err = nullptr, *cand_done = nullptr;
  uint64_t* cand_body = nullptr;
  uint8_t *cand_esc = nullptr, *stripe_state = nullptr;
  uint32_t* tile_state = nullptr;
  uint32_t stripe_cap = 0, tile_cap = 0;
  uint32_t epoch = 0;                     // pass number on this device tree (m2k_scan increments it)
  uint32_t* special = nullptr;            // [0] count, then ids of entries with a special name (compacted by k_name_features)
  uint32_t special_cap = 0;
  // One allocation holds every array a pass starts from a known value, so that a pass begins with two
  // memsets instead of nine: a zero region (dir_direct, dir_content, dir_subtree, file_flags, the
  // isDockerFile counters, special[0]) followed by the list part of `special`, then an all-ones
  // region (dir_first_war, dir_main, dir_main_sub: "none yet" for the MIN reductions).
  uint8_t* early = nullptr;               // dir_opts | dir_state | dir_ignore: one allocation, one copy to the host
  size_t early_bytes = 0;
  uint8_t* slab = nullptr;
  size_t slab_zero_bytes = 0, slab_ff_off = 0, slab_ff_bytes = 0, slab_bytes = 0;
  // behind the all-ones region, never cleared by a pass: the exchange mailbox and the gathered piece summaries
  Mailbox* mailbox = nullptr;
  PieceSummary* all_pieces = nullptr;     // [2 * kMaxRanks]
  PeerLink link;
  void* ign_blob = nullptr;               // ignore table + folded runes of the last finalize (grow-only)
  size_t ign_blob_cap = 0;
  uint32_t* entry_feat = nullptr;         // [n_entries]
  uint32_t* dir_direct = nullptr;         // [n_dirs]
  uint32_t* dir_first_war = nullptr;      // [n_dirs]
  uint32_t* dir_content = nullptr;        // [n_dirs] partial: content bits of this shard
  uint32_t* dir_main = nullptr;           // [n_dirs] partial: lowest main-file id of this shard
  uint32_t* dir_subtree = nullptr;        // [n_dirs]
  uint32_t* dir_main_sub = nullptr;       // [n_dirs]
  uint32_t* dir_opts = nullptr;           // [n_dirs]
  uint8_t* dir_ignore = nullptr;          // [n_dirs]
  uint8_t* dir_state = nullptr;           // [n_dirs]
  uint8_t* dir_descend = nullptr;         // [n_dirs]
  uint8_t* file_flags = nullptr;          // [n_files] (only the shard's range is written)
  bool scanned = false;
};

struct Ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[12] = {};
  // second stream for the small copies that would otherwise sit between kernels of `stream` (name
  // features to the host, the ignore table to the device, early result arrays), and its ordering events
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_names = nullptr, ev_feat = nullptr, ev_blob = nullptr, ev_fork = nullptr, ev_join = nullptr;
  // third stream, highest priority: the directory-level phases (name features up the tree, the 18
  // predicates, the ignore join and the walk) depend on names only, so they run while the byte scan
  // still owns the main stream; their blocks are placed as scan blocks retire
  cudaStream_t side_stream = nullptr;
  cudaEvent_t ev_side0 = nullptr, ev_side1 = nullptr;  // timing on: the span of those kernels
  cudaEvent_t ev_early = nullptr;                      // ... and their results are on the host
  cudaEvent_t ev_idle = nullptr;                       // blocking-sync event for ranks that only wait (m2k_exchange_wait)
  cudaEvent_t ev_pass[2] = {nullptr, nullptr};         // blocking-sync: an attached rank's pass p has been delivered (p & 1)
  // two pinned blocks through which bytes held in pageable host memory are streamed to the device
  void* stage[2] = {nullptr, nullptr};
  cudaEvent_t stage_ev[2] = {nullptr, nullptr};
  // the two allocations that grow with the shard -- its bytes and the Dockerfile stripe states -- are kept by the
  // context when a device tree is released and handed to the next one that fits (grow-only): a plugin that
  // detects tree after tree does not pay cudaMalloc / cudaFree of tens of GB per call (m2k_ctx_trim drops them)
  struct Arena { void* p = nullptr; size_t cap = 0; bool in_use = false; } arena[2];
  // page-locked host blocks (result arrays, name features, the ignore table's staging) are pooled the same way:
  // page-locking and unlocking per call costs milliseconds, and far more once the process holds a lot of
  // pinned memory (measured: ~200 ms per m2k_detect next to a 99 GB pinned tree)
  struct PinnedBlock { void* p; size_t cap; bool in_use; };
  std::vector<PinnedBlock> pinned_pool;
  float phase_ms[8] = {};  // names, scan, combine, dockerfile, finalize kernels, host ignore, host verify, d2h
  int sm_count = 0;
  uint64_t launches = 0;
  float last_scan_ms = 0.f;
  std::string err;
  uint32_t *d_fold_from = nullptr, *d_fold_to = nullptr;
  NameSig* d_sigs = nullptr;
  int n_sigs = 0;
};

// ---- kernel launchers (kernels.cu) ----
void launch_name_features(Ctx* c, DevTree* t);
void launch_scan_bytes(Ctx* c, DevTree* t, ChunkRec* chunk_rec, uint32_t* ticket);
void launch_file_phase(Ctx* c, DevTree* t, const ChunkRec* chunk_rec);
void launch_dockerfile(Ctx* c, DevTree* t);
void launch_content_up(Ctx* c, DevTree* t, cudaStream_t st);
void launch_walk(Ctx* c, DevTree* t, uint32_t sh_flavor, const IgnoreSlot* table, uint32_t table_mask, const uint32_t* runes,
                 const uint32_t* prefix_runes, uint32_t prefix_len, uint64_t depth1_state, uint8_t root_ignore, cudaStream_t st);
void fold_runes(const std::string& s, std::vector<uint32_t>* out);
std::string fast_dir_path(const m2k_tree* t, uint32_t d, const std::string& clean_root);
void launch_merge_partials(Ctx* c, DevTree* t, const uint32_t* all_content, const uint32_t* all_main, uint32_t n_ranks);
void launch_merge_pieces(Ctx* c, DevTree* t, const PieceSummary* all_pieces, uint32_t n_pieces);
void launch_settle_file(Ctx* c, DevTree* t, uint32_t f, uint32_t flags);
void launch_synth_fill(Ctx* c, DevTree* t, uint64_t seed);
// exchange (kernels.cu): flag writes / waits at system scope, this rank's flags + pieces into the root's memory
void launch_signal(Ctx* c, uint32_t* flag, uint32_t value, cudaStream_t st);
void launch_wait_ge(Ctx* c, const uint32_t* flags, uint32_t n_flags, uint32_t value, uint32_t* abort, cudaStream_t st);
void launch_contribute(Ctx* c, DevTree* t, cudaStream_t st);
int preload_kernels();  // 0 or a cudaError_t
int read_violations(unsigned long long* out);  // checked builds: the bounds-check counters (and clears them)
int checked_build();
void launch_fill_u32(Ctx* c, uint32_t* p, uint32_t v, size_t n);

void set_global_error(const char* msg);  // api.cu: what m2k_global_error() returns on this thread

// ---- host helpers (hostpath.cc) ----
// cleaned paths packed into one character array: entry i is chars[off[i] .. off[i+1])
struct PathArena {
  std::vector<char> chars;
  std::vector<uint32_t> off;
  std::vector<uint8_t> flag;
  bool missing = false;
  size_t size() const { return flag.size(); }
};
size_t path_clean_buf(const char* p, size_t n, char* out);
void parse_ignore_file_arena(const std::string& dir, const uint8_t* data, uint64_t len, PathArena* out);
std::string path_clean(const std::string& p);
std::string path_join(const std::string& a, const std::string& b);
std::string path_dir(const std::string& p);
bool equal_fold(const std::string& a, const std::string& b);
uint64_t hash_fold(uint64_t h, const std::string& s);
void parse_ignore_file(const std::string& file_path, const uint8_t* data, uint64_t len,
                       std::vector<std::string>* ignore_dirs, std::vector<std::string>* ignore_contents);
std::string tree_dir_path(const m2k_tree* t, uint32_t d);
void sort_walk_order(const m2k_tree* t, std::vector<uint32_t>* ents);    // a set of entry ids into filepath.Walk order

// ---- tree ownership (packer.cc / synth.cc) ----
struct TreeOwner {
  bool pinned = false;
  std::string root;
  std::vector<uint8_t> synth_class;    // synthetic trees only
  std::vector<uint64_t> synth_inject;
  uint64_t synth_seed = 0;
  bool synthetic = false;
  // files whose bytes are literal rather than procedural (.m2kignore files of synthetic trees)
  std::vector<uint32_t> literal_file;
  std::vector<std::string> literal_data;
};
m2k_tree* tree_alloc(uint32_t n_dirs, uint32_t n_entries, uint32_t n_files, uint64_t n_name_bytes, uint64_t n_bytes,
                     bool with_bytes, bool pinned, const char* root);
void tree_release(m2k_tree* t);
int pack_tree_impl(bool pinned, const char* root_path, int read_content, m2k_tree** out, std::string* err);
// the packer in steps, for callers that stream the bytes instead of holding them (m2k_pack_detect):
// the walk + numbering + byte layout (a tree with or without a bytes blob), then any byte range on demand
struct PackPlan { struct Impl; Impl* p; PackPlan(); ~PackPlan(); PackPlan(const PackPlan&) = delete; PackPlan& operator=(const PackPlan&) = delete; };
int pack_plan(const char* root_path, int read_content, bool pinned, bool with_bytes, PackPlan* plan, m2k_tree** out, std::string* err);
uint64_t pack_plan_bytes(const PackPlan* plan);
bool pack_read_range(const PackPlan* plan, uint64_t lo, uint64_t hi, uint8_t* dst);
void pack_keep_rule_files(const PackPlan* plan, m2k_tree* t);
int synth_tree_impl(bool pinned, const m2k_synth_params* p, int host_bytes, m2k_tree** out, std::string* err);
void synth_fill_host(const m2k_tree* tree, uint32_t file_begin, uint32_t file_end, uint8_t* dst, int n_threads);

}  // namespace m2k
