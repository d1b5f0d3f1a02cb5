// C ABI of libm2kdetect (include/m2kdetect.h): contexts, uploads, the kernel pipeline, results.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <unordered_map>

#include "casefold_table.h"
#include "gostr.h"
#include "m2k_internal.h"

using namespace m2k;

struct m2k_ctx { Ctx c; };
constexpr uint32_t kSpecialCap = 1u << 16;  // entries with a special name (.m2kignore, user detect scripts) listed by the device

// what m2k_prepare / m2k_finalize derive from the .m2kignore files (host side of a2 / a4)
struct Folded { uint64_t key; uint32_t off, len, flag; };  // one ignore entry: hash of its folded runes, where they are

// Results and device trees may outlive their context (a garbage-collected host language releases in any order):
// what they hold from a context's pools goes back to the pool only while that context is alive, and is
// freed directly otherwise.
static std::mutex g_live_mu;
static std::vector<Ctx*> g_live_ctx;
static bool ctx_alive(Ctx* C) {  // caller holds g_live_mu
  return C && std::find(g_live_ctx.begin(), g_live_ctx.end(), C) != g_live_ctx.end();
}

// pinned host blocks from the context's pool (Ctx::pinned_pool); a NULL context means plain cudaHostAlloc
static void* pinned_take(Ctx* C, size_t bytes) {
  if (!bytes) bytes = 16;
  if (C) {  // the pool's book-keeping under the registry's lock: a result may be released from another thread meanwhile
    std::lock_guard<std::mutex> lk(g_live_mu);
    Ctx::PinnedBlock* best = nullptr;
    for (auto& b : C->pinned_pool)
      if (!b.in_use && b.cap >= bytes && (!best || b.cap < best->cap)) best = &b;
    if (best && best->cap <= 4 * bytes + (1u << 20)) { best->in_use = true; return best->p; }
  }
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  if (C) { std::lock_guard<std::mutex> lk(g_live_mu); C->pinned_pool.push_back({p, bytes, true}); }
  return p;
}
static void pinned_give(Ctx* C, void* p) {
  if (!p) return;
  {
    std::lock_guard<std::mutex> lk(g_live_mu);
    if (ctx_alive(C))
      for (auto& b : C->pinned_pool)
        if (b.p == p) { b.in_use = false; return; }
  }
  cudaFreeHost(p);  // not pooled, or the pool's context is gone (it left its lent blocks to their holders)
}

// grow-only host buffer the ignore table is built in.  Pinned in the product: a copy from pageable
// memory would make m2k_prepare wait for the byte scan queued ahead of it.
struct Staging {
  void* p = nullptr;
  size_t cap = 0;
  bool pinned = true;
  Ctx* pool = nullptr;  // pinned blocks come from / go back to this context's pool
  // room for `bytes`; the first `keep` bytes survive a reallocation
  bool reserve(size_t bytes, size_t keep) {
    if (bytes <= cap) return true;
    const size_t want = bytes * 2;
    void* q = nullptr;
    if (pinned) { if (!(q = pinned_take(pool, want))) return false; }
    else if (!(q = malloc(want))) return false;
    if (keep) memcpy(q, p, keep);
    release();
    p = q; cap = want;
    return true;
  }
  void release() {
    if (p) { if (pinned) pinned_give(pool, p); else free(p); }
    p = nullptr; cap = 0;
  }
};

struct IgnorePrep {
  bool ready = false;
  // cleaned entries in the reference's list order; the last arena = preContainerized.  The vectors
  // below keep their capacity from pass to pass (fresh megabytes cost more in page faults than the work)
  std::vector<PathArena> arenas;
  std::vector<uint32_t> ign_entries;
  std::vector<Folded> folded;
  uint32_t table_mask = M2K_NONE, prefix_len = 0;
  uint64_t depth1_state = 0;
  uint8_t root_ignore = 0;
  // layout of the staged blob: folded runes, the folded depth-1 prefix, then (16-byte aligned) the table
  size_t blob_bytes = 0, off_prefix = 0, off_table = 0;
  const IgnoreSlot* d_table = nullptr;
  const uint32_t *d_runes = nullptr, *d_prefix = nullptr;
  float ms_host = 0.f;
  uint32_t sh_flavor = M2K_SH_BASH;     // the /bin/sh flavour the early directory phases were run for
  double ms_part[5] = {0, 0, 0, 0, 0};  // order, parse, fold, table, lists
  void reset() {
    ready = false; table_mask = M2K_NONE; prefix_len = 0; depth1_state = 0; root_ignore = 0;
    blob_bytes = off_prefix = off_table = 0;
    d_table = nullptr; d_runes = d_prefix = nullptr; ms_host = 0.f;
  }
};

struct ListBuf { char* p = nullptr; size_t cap = 0; };  // grow-only block one ignore list of a result lives in

struct m2k_devtree {
  DevTree t;
  ListBuf lists[2];                  // ignore_dirs / ignore_contents of the pending or last result
  float ms_h2d = 0.f;
  uint32_t* h_entry_feat = nullptr;  // pinned
  uint32_t* h_special = nullptr;     // pinned: [0] count, then entry ids
  Staging stage;                     // pinned staging of the ignore table (grow-only)
  Ctx* ctx = nullptr;                // the context whose pinned pool the host buffers below belong to
  m2k_result* pending = nullptr;     // result shell made by m2k_prepare
  uint8_t* h_results = nullptr;      // pinned: every per-directory / per-file result array (see result_layout)
  bool content = true;
  IgnorePrep prep;
};

static thread_local std::string g_err;

static int fail(Ctx* c, int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (c) c->err = buf;
  g_err = buf;
  return code;
}

namespace m2k {
void set_global_error(const char* msg) { g_err = msg ? msg : ""; }
}  // namespace m2k

#define CU(call)                                                                                              \
  do {                                                                                                        \
    cudaError_t e_ = (call);                                                                                  \
    if (e_ != cudaSuccess) return fail(C, M2K_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---- name signature table -------------------------------------------------------------------
static NameSig make_sig(const char* pat, uint32_t kind, uint32_t feature, uint32_t need) {
  NameSig g;
  memset(&g, 0, sizeof g);
  size_t n = strlen(pat);
  g.len = (uint32_t)n; g.kind = kind; g.feature = feature; g.need = need;
  for (size_t i = 0; i < n && i < 16; i++) {
    uint64_t ch = (uint8_t)pat[n - 1 - i];
    if (i < 8) { g.lo |= ch << (8 * i); g.mask_lo |= 0xFFull << (8 * i); }
    else { g.hi |= ch << (8 * (i - 8)); g.mask_hi |= 0xFFull << (8 * (i - 8)); }
  }
  return g;
}

static std::vector<NameSig> default_sigs() {
  std::vector<NameSig> s;
  // [ -f D/name ]  (exact name, regular file after following symlinks)
  s.push_back(make_sig("Pipfile", 0, M2K_F_PIPFILE, 1));
  s.push_back(make_sig("build.xml", 0, M2K_F_BUILD_XML, 1));
  s.push_back(make_sig("build.gradle", 0, M2K_F_BUILD_GRADLE, 1));
  s.push_back(make_sig("pom.xml", 0, M2K_F_POM_XML, 1));
  s.push_back(make_sig("package.json", 0, M2K_F_PACKAGE_JSON, 1));
  s.push_back(make_sig("requirements.txt", 0, M2K_F_REQUIREMENTS, 1));
  s.push_back(make_sig("setup.py", 0, M2K_F_SETUP_PY, 1));
  s.push_back(make_sig("environment.yml", 0, M2K_F_ENVIRONMENT_YML, 1));
  s.push_back(make_sig("Gemfile", 0, M2K_F_GEMFILE, 1));
  s.push_back(make_sig("go.mod", 0, M2K_F_GO_MOD, 1));
  // find -name "*.ext"  (any entry type, dotfiles included)
  s.push_back(make_sig(".go", 1, M2K_R_GO, 0));
  s.push_back(make_sig(".php", 1, M2K_R_PHP, 0));
  s.push_back(make_sig(".java", 1, M2K_R_JAVA, 0));
  // "$1/"*.war  (shell glob: no dotfiles)
  s.push_back(make_sig(".war", 1, M2K_N_WAR, 3));
  // GetFilesByName targets (non-directories)
  s.push_back(make_sig(".m2kignore", 0, M2K_N_IGNOREFILE, 2));
  s.push_back(make_sig("m2kdfdetect.sh", 0, M2K_N_USER_DF, 2));
  s.push_back(make_sig("m2ks2idetect.sh", 0, M2K_N_USER_S2I, 2));
  // GetFilesByExt(.yml, .yaml) (non-directories)
  s.push_back(make_sig(".yml", 1, M2K_N_YAML, 2));
  s.push_back(make_sig(".yaml", 1, M2K_N_YAML, 2));
  return s;
}

// ---- lifecycle ------------------------------------------------------------------------------
extern "C" int m2k_abi_version(void) { return M2K_ABI_VERSION; }
extern "C" const char* m2k_global_error(void) { return g_err.c_str(); }
extern "C" const char* m2k_last_error(const m2k_ctx* ctx) { return ctx ? ctx->c.err.c_str() : g_err.c_str(); }

extern "C" int m2k_ctx_create(int device, m2k_ctx** out) {
  if (!out) return fail(nullptr, M2K_ERR_ARG, "m2k_ctx_create: out is NULL");
  *out = nullptr;
  Ctx* C = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(nullptr, M2K_ERR_CUDA, "no CUDA device (%s): libm2kdetect has no CPU path", e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
  }
  if (device < 0 || device >= n) return fail(nullptr, M2K_ERR_ARG, "device %d out of range (0..%d)", device, n - 1);
  m2k_ctx* ctx = new m2k_ctx();
  C = &ctx->c;
  C->device = device;
  auto bail = [&](int rc) { std::string keep = C->err; delete ctx; g_err = keep; return rc; };
#define CUX(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail(C, M2K_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); return bail(M2K_ERR_CUDA); } } while (0)
  CUX(cudaSetDevice(device));
  CUX(cudaStreamCreateWithFlags(&C->stream, cudaStreamNonBlocking));
  for (auto& ev : C->ev) CUX(cudaEventCreate(&ev));
  CUX(cudaStreamCreateWithFlags(&C->copy_stream, cudaStreamNonBlocking));
  for (cudaEvent_t* ev : {&C->ev_names, &C->ev_feat, &C->ev_blob, &C->ev_fork, &C->ev_join, &C->ev_early}) CUX(cudaEventCreateWithFlags(ev, cudaEventDisableTiming));
  {
    int least = 0, greatest = 0;
    CUX(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    CUX(cudaStreamCreateWithPriority(&C->side_stream, cudaStreamNonBlocking, greatest));
    CUX(cudaEventCreate(&C->ev_side0));
    CUX(cudaEventCreate(&C->ev_side1));
  }
  CUX(cudaDeviceGetAttribute(&C->sm_count, cudaDevAttrMultiProcessorCount, device));
  CUX(cudaMalloc(&C->d_fold_from, sizeof(m2k_casefold_from)));
  CUX(cudaMalloc(&C->d_fold_to, sizeof(m2k_casefold_to)));
  CUX(cudaMemcpy(C->d_fold_from, m2k_casefold_from, sizeof(m2k_casefold_from), cudaMemcpyHostToDevice));
  CUX(cudaMemcpy(C->d_fold_to, m2k_casefold_to, sizeof(m2k_casefold_to), cudaMemcpyHostToDevice));
  std::vector<NameSig> sigs = default_sigs();
  C->n_sigs = (int)sigs.size();
  CUX(cudaMalloc(&C->d_sigs, sizeof(NameSig) * sigs.size()));
  CUX(cudaMemcpy(C->d_sigs, sigs.data(), sizeof(NameSig) * sigs.size(), cudaMemcpyHostToDevice));
#undef CUX
  {
    std::lock_guard<std::mutex> lk(g_live_mu);
    g_live_ctx.push_back(C);
  }
  *out = ctx;
  return M2K_OK;
}

extern "C" void m2k_ctx_destroy(m2k_ctx* ctx) {
  if (!ctx) return;
  Ctx* C = &ctx->c;
  cudaSetDevice(C->device);
  if (C->stream) cudaStreamSynchronize(C->stream);
  if (C->copy_stream) cudaStreamSynchronize(C->copy_stream);
  if (C->side_stream) cudaStreamSynchronize(C->side_stream);
  cudaFree(C->d_fold_from); cudaFree(C->d_fold_to); cudaFree(C->d_sigs);
  {
    std::lock_guard<std::mutex> lk(g_live_mu);
    g_live_ctx.erase(std::remove(g_live_ctx.begin(), g_live_ctx.end(), C), g_live_ctx.end());
  }
  // what is lent out stays with its holder (a result or device tree released later frees it itself)
  for (auto& A : C->arena) if (A.p && !A.in_use) cudaFree(A.p);
  for (auto& b : C->pinned_pool) if (!b.in_use) cudaFreeHost(b.p);
  for (auto& ev : C->ev) if (ev) cudaEventDestroy(ev);
  for (cudaEvent_t ev : {C->ev_names, C->ev_feat, C->ev_blob, C->ev_fork, C->ev_join, C->ev_early, C->ev_side0, C->ev_side1}) if (ev) cudaEventDestroy(ev);
  if (C->ev_idle) cudaEventDestroy(C->ev_idle);
  for (cudaEvent_t ev : C->ev_pass) if (ev) cudaEventDestroy(ev);
  if (C->side_stream) cudaStreamDestroy(C->side_stream);
  for (int k = 0; k < 2; k++) { if (C->stage[k]) cudaFreeHost(C->stage[k]); if (C->stage_ev[k]) cudaEventDestroy(C->stage_ev[k]); }
  if (C->stream) cudaStreamDestroy(C->stream);
  if (C->copy_stream) cudaStreamDestroy(C->copy_stream);
  delete ctx;
}

extern "C" int m2k_device_info(m2k_ctx* ctx, int* sm_count, uint64_t* hbm_bytes, char* name, size_t name_cap) {
  if (!ctx) return fail(nullptr, M2K_ERR_ARG, "ctx is NULL");
  Ctx* C = &ctx->c;
  cudaDeviceProp p;
  CU(cudaGetDeviceProperties(&p, C->device));
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (hbm_bytes) *hbm_bytes = p.totalGlobalMem;
  if (name && name_cap) { strncpy(name, p.name, name_cap - 1); name[name_cap - 1] = 0; }
  return M2K_OK;
}

// diagnostic: the counters of a checked build (kernels.cu M2K_CHECKED): out[0..8), cleared by the call.
// Returns 1 when the library was built with the checks, 0 when not (the counters then stay zero).
extern "C" int m2k_debug_violations(m2k_ctx* ctx, uint64_t* out) {
  if (!ctx || !out) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  CU(cudaDeviceSynchronize());
  unsigned long long v[8];
  if (int e = read_violations(v)) return fail(C, M2K_ERR_CUDA, "reading the check counters: %s", cudaGetErrorString((cudaError_t)e));
  for (int i = 0; i < 8; i++) out[i] = v[i];
  return checked_build();
}

extern "C" int m2k_ctx_trim(m2k_ctx* ctx) {
  if (!ctx) return fail(nullptr, M2K_ERR_ARG, "ctx is NULL");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  std::vector<void*> dev, host;  // taken out of the books under the lock, given back to the driver outside it
  {
    std::lock_guard<std::mutex> lk(g_live_mu);
    for (auto& A : C->arena) if (A.p && !A.in_use) { dev.push_back(A.p); A.p = nullptr; A.cap = 0; }
    for (size_t i = 0; i < C->pinned_pool.size();) {
      if (!C->pinned_pool[i].in_use) { host.push_back(C->pinned_pool[i].p); C->pinned_pool.erase(C->pinned_pool.begin() + i); }
      else i++;
    }
  }
  for (void* p : dev) cudaFree(p);
  for (void* p : host) cudaFreeHost(p);
  return M2K_OK;
}

extern "C" void* m2k_stream(m2k_ctx* ctx) { return ctx ? (void*)ctx->c.stream : nullptr; }
extern "C" uint64_t m2k_launch_count(const m2k_ctx* ctx) { return ctx ? ctx->c.launches : 0; }
extern "C" int m2k_phase_ms(m2k_ctx* ctx, float* out, int cap) {
  if (!ctx || !out) return fail(nullptr, M2K_ERR_ARG, "bad argument");
  for (int i = 0; i < cap && i < 8; i++) out[i] = ctx->c.phase_ms[i];
  return M2K_OK;
}
extern "C" int m2k_scan_kernel_ms(m2k_ctx* ctx, float* ms) {
  if (!ctx || !ms) return fail(nullptr, M2K_ERR_ARG, "bad argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  CU(cudaStreamSynchronize(C->stream));
  if (cudaEventElapsedTime(&C->last_scan_ms, C->ev[2], C->ev[3]) != cudaSuccess) cudaGetLastError();
  *ms = C->last_scan_ms;
  return M2K_OK;
}

// ---- host trees -----------------------------------------------------------------------------
extern "C" int m2k_pack_tree(m2k_ctx* ctx, const char* root_path, int read_content, m2k_tree** out) {
  if (!root_path || !out) return fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_pack_tree: NULL argument");
  if (ctx) cudaSetDevice(ctx->c.device);
  std::string err;
  int rc = pack_tree_impl(ctx != nullptr, root_path, read_content, out, &err);
  if (rc != M2K_OK) return fail(ctx ? &ctx->c : nullptr, rc, "%s", err.c_str());
  return rc;
}

extern "C" int m2k_synth_tree(m2k_ctx* ctx, const m2k_synth_params* p, int host_bytes, m2k_tree** out) {
  if (!p || !out) return fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_synth_tree: NULL argument");
  if (ctx) cudaSetDevice(ctx->c.device);
  std::string err;
  int rc = synth_tree_impl(ctx != nullptr, p, host_bytes, out, &err);
  if (rc != M2K_OK) return fail(ctx ? &ctx->c : nullptr, rc, "%s", err.c_str());
  return rc;
}

extern "C" int m2k_synth_fill_host(const m2k_tree* tree, uint64_t seed, uint32_t file_begin, uint32_t file_end, uint8_t* dst,
                                   int n_threads) {
  if (!tree || !dst || file_begin > file_end || file_end > tree->n_files) return fail(nullptr, M2K_ERR_ARG, "m2k_synth_fill_host: bad argument");
  const TreeOwner* o = static_cast<const TreeOwner*>(tree->owner);
  if (!o || !o->synthetic) return fail(nullptr, M2K_ERR_ARG, "not a synthetic tree");
  if (seed != o->synth_seed) return fail(nullptr, M2K_ERR_ARG, "seed differs from the tree's seed");
  synth_fill_host(tree, file_begin, file_end, dst, n_threads);
  return M2K_OK;
}

// pinned host arena for callers that stage file bytes themselves (never hand
// garbage-collected memory to asynchronous copies)
extern "C" void* m2k_pinned_alloc(m2k_ctx* ctx, size_t bytes) {
  if (!ctx || !bytes) { fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_pinned_alloc: bad argument"); return nullptr; }
  cudaSetDevice(ctx->c.device);
  void* p = nullptr;
  cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocPortable);
  if (e != cudaSuccess) { cudaGetLastError(); fail(&ctx->c, M2K_ERR_NOMEM, "cudaHostAlloc(%zu bytes): %s", bytes, cudaGetErrorString(e)); return nullptr; }
  return p;
}
extern "C" void m2k_pinned_free(m2k_ctx* ctx, void* p) {
  if (ctx) cudaSetDevice(ctx->c.device);
  if (p) cudaFreeHost(p);
}

extern "C" void m2k_tree_free(m2k_ctx* ctx, m2k_tree* tree) {
  if (ctx) cudaSetDevice(ctx->c.device);
  tree_release(tree);
}

// ---- device trees ---------------------------------------------------------------------------
static void devtree_release(DevTree* t) {
  if (t->link.ipc_base) { cudaIpcCloseMemHandle(t->link.ipc_base); t->link.ipc_base = nullptr; }
  // (dir_direct, dir_content, dir_subtree, file_flags, df_counters, special, dir_first_war, dir_main and
  // dir_main_sub live inside the slab)
  if (t->arena_ctx) {  // hand the big buffers back to the context instead of freeing them -- if it is still there
    std::lock_guard<std::mutex> lk(g_live_mu);
    if (ctx_alive(t->arena_ctx)) {
      if (t->from_arena & 1u) { t->arena_ctx->arena[0].in_use = false; t->bytes = nullptr; }
      if (t->from_arena & 2u) { t->arena_ctx->arena[1].in_use = false; t->stripe_state = nullptr; }
    }
    t->from_arena = 0;
    t->arena_ctx = nullptr;
  }
  void* ptrs[] = {t->dir_parent, t->dir_entry, t->dir_depth, t->dir_flags, t->entry_dir, t->entry_name_off,
                  t->entry_child, t->entry_type, t->names, t->file_entry, t->file_off, t->file_len, t->chunk_first, t->chunk_file, t->bytes,
                  t->synth_class, t->synth_inject, t->chunk_rec, t->entry_feat, t->slab, t->early,
                  t->dir_descend, t->head_list, t->cand_file, t->cand_stripe0, t->cand_err, t->cand_done, t->cand_body,
                  t->cand_esc, t->stripe_state, t->tile_state, t->ign_blob, t->pieces};
  for (void* p : ptrs) if (p) cudaFree(p);
}

// pinned host buffers and the object itself (device memory is released separately)
static void free_devtree_shell(m2k_devtree* dt) {
  pinned_give(dt->ctx, dt->h_entry_feat);
  pinned_give(dt->ctx, dt->h_special);
  dt->stage.release();
  free(dt->lists[0].p); free(dt->lists[1].p);
  pinned_give(dt->ctx, dt->h_results);
  if (dt->pending) { m2k_result* p = dt->pending; dt->pending = nullptr; p->owner = nullptr; m2k_result_free(p); }
  delete dt;
}

extern "C" void m2k_devtree_free(m2k_ctx* ctx, m2k_devtree* dt) {
  if (!dt) return;
  if (ctx) { cudaSetDevice(ctx->c.device); cudaStreamSynchronize(ctx->c.stream); cudaStreamSynchronize(ctx->c.copy_stream); cudaStreamSynchronize(ctx->c.side_stream); }
  devtree_release(&dt->t);
  free_devtree_shell(dt);
}

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// `bytes` bytes from the context's arena k when it is free (growing it when too small), else a plain allocation
static cudaError_t arena_alloc(Ctx* C, DevTree* t, int k, void** out, size_t bytes) {
  Ctx::Arena& A = C->arena[k];
  if (!bytes) bytes = 1;
  bool busy;
  {  // claim the arena under the registry's lock (a device tree may hand it back from another thread)
    std::lock_guard<std::mutex> lk(g_live_mu);
    busy = A.in_use;
    if (!busy) A.in_use = true;
  }
  if (busy) return cudaMalloc(out, bytes);
  if (A.cap < bytes) {
    if (A.p) cudaFree(A.p);
    A.p = nullptr; A.cap = 0;
    cudaError_t e = cudaMalloc(&A.p, bytes);
    if (e != cudaSuccess) { std::lock_guard<std::mutex> lk(g_live_mu); A.in_use = false; return e; }
    A.cap = bytes;
  }
  t->arena_ctx = C;
  t->from_arena |= 1u << k;
  *out = A.p;
  return cudaSuccess;
}

template <typename T>
static cudaError_t dev_alloc(T** p, size_t n) { return cudaMalloc((void**)p, (n ? n : 1) * sizeof(T)); }

static void parallel_memcpy(void* dst, const void* src, size_t n);  // below, on the host thread team

// Host-to-device copy of `n` bytes on stream s.  From pinned (or registered) memory it is one DMA.
// From pageable memory the driver would stage the copy itself, one thread, ~10 GB/s; instead the
// bytes go through two 32 MiB pinned blocks owned by the context -- a team of host threads fills one
// while the DMA engine drains the other -- which needs no page-locking of the caller's buffer (that
// costs more than reading the files did: 0.6 s for 2 GB).
static constexpr size_t kStageBytes = 32u << 20;
static int ensure_stage(Ctx* C) {
  for (int k = 0; k < 2; k++) {
    if (!C->stage[k]) { CU(cudaHostAlloc(&C->stage[k], kStageBytes, cudaHostAllocDefault)); CU(cudaEventCreateWithFlags(&C->stage_ev[k], cudaEventDisableTiming)); }
    else CU(cudaEventSynchronize(C->stage_ev[k]));
  }
  return M2K_OK;
}
static int copy_to_device(Ctx* C, cudaStream_t s, void* dst, const void* src, size_t n) {
  if (!n) return M2K_OK;
  cudaPointerAttributes attr;
  bool pageable = false;
  if (cudaPointerGetAttributes(&attr, src) == cudaSuccess) pageable = attr.type == cudaMemoryTypeUnregistered;
  else cudaGetLastError();
  if (!pageable || n < (4u << 20)) { CU(cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, s)); return M2K_OK; }
  if (int rc = ensure_stage(C)) return rc;
  size_t off = 0;
  for (int k = 0; off < n; k ^= 1) {
    const size_t len = std::min(kStageBytes, n - off);
    CU(cudaEventSynchronize(C->stage_ev[k]));  // the block's previous DMA is done (a never-recorded event is complete)
    parallel_memcpy(C->stage[k], static_cast<const char*>(src) + off, len);
    CU(cudaMemcpyAsync(static_cast<char*>(dst) + off, C->stage[k], len, cudaMemcpyHostToDevice, s));
    CU(cudaEventRecord(C->stage_ev[k], s));
    off += len;
  }
  return M2K_OK;
}

static int copy_shard(Ctx* C, m2k_devtree* dt, const m2k_tree* tree);

static inline uint64_t file_chunks(const m2k_tree* tree, uint32_t f) { return (tree->file_len[f] + kChunkBytes - 1) / kChunkBytes; }

// the shard: files [file_begin, file_end), of which the first starts `first_skip` chunks in and the
// last ends at chunk `last_end` of that file (M2K_NONE: at its end)
static int upload_impl(m2k_ctx* ctx, const m2k_tree* tree, uint32_t file_begin, uint32_t file_end, uint32_t first_skip, uint32_t last_end,
                       m2k_devtree** out) {
  Ctx* C = &ctx->c;
  *out = nullptr;
  const double t_setup = now_ms();
  CU(cudaSetDevice(C->device));
  m2k_devtree* dt = new m2k_devtree();
  DevTree* t = &dt->t;
  struct Guard { m2k_devtree* d; bool keep = false; ~Guard() { if (!keep) { devtree_release(&d->t); free_devtree_shell(d); } } } guard{dt};
  t->host = tree;
  t->n_dirs = tree->n_dirs; t->n_entries = tree->n_entries; t->n_files = tree->n_files; t->max_depth = tree->max_depth;
  t->file_begin = file_begin; t->file_end = file_end; t->first_skip = first_skip; t->last_end = last_end;
  const uint32_t nd = t->n_dirs, ne = t->n_entries, nf = t->n_files, ns = file_end - file_begin;
  // chunk prefix over the shard
  std::vector<uint32_t> chunk_first((size_t)ns + 1);
  uint64_t nchunks = 0, content = 0;
  for (uint32_t i = 0; i < ns; i++) {
    chunk_first[i] = (uint32_t)nchunks;
    const uint64_t len = tree->file_len[file_begin + i];
    const uint64_t c_lo = i == 0 ? first_skip : 0;
    const uint64_t c_hi = (i + 1 == ns && last_end != M2K_NONE) ? last_end : file_chunks(tree, file_begin + i);
    nchunks += c_hi - c_lo;
    const uint64_t b_lo = c_lo * kChunkBytes, b_hi = std::min<uint64_t>(len, c_hi * kChunkBytes);
    content += b_hi - b_lo;
  }
  chunk_first[ns] = (uint32_t)nchunks;
  if (nchunks > 0xFFFFFFF0ull) return fail(C, M2K_ERR_LIMIT, "shard has too many 64 KiB chunks (%llu)", (unsigned long long)nchunks);
  t->n_chunks = (uint32_t)nchunks;
  t->shard_content = content;
  // resident bytes: the shard's chunks, plus 16 bytes on the far side of a cut through a file (a
  // "__main__" may straddle the cut), plus 16 zero bytes at the very end
  uint64_t r0 = 0, r1 = 0;
  if (ns) {
    r0 = tree->file_off[file_begin] + (uint64_t)first_skip * kChunkBytes - (first_skip ? 16 : 0);
    const uint32_t fl = file_end - 1;
    r1 = last_end != M2K_NONE ? tree->file_off[fl] + (uint64_t)last_end * kChunkBytes + 16
                              : tree->file_off[fl] + ((tree->file_len[fl] + 15) & ~15ull);
  }
  t->shard_byte0 = r0;
  t->shard_bytes = r1 - r0 + 16;

  CU(dev_alloc(&t->dir_parent, nd)); CU(dev_alloc(&t->dir_entry, nd)); CU(dev_alloc(&t->dir_depth, nd));
  CU(dev_alloc(&t->dir_flags, nd));
  CU(dev_alloc(&t->entry_dir, ne)); CU(dev_alloc(&t->entry_name_off, (size_t)ne + 1)); CU(dev_alloc(&t->entry_child, ne));
  CU(dev_alloc(&t->entry_type, ne)); CU(dev_alloc(&t->names, (size_t)tree->n_name_bytes));
  CU(dev_alloc(&t->file_entry, nf)); CU(dev_alloc(&t->file_off, nf)); CU(dev_alloc(&t->file_len, nf));
  CU(dev_alloc(&t->chunk_first, (size_t)ns + 1)); CU(dev_alloc(&t->chunk_file, t->n_chunks));
  CU(arena_alloc(C, t, 0, (void**)&t->bytes, (size_t)t->shard_bytes));
  CU(dev_alloc(&t->chunk_rec, t->n_chunks)); CU(dev_alloc(&t->entry_feat, ne)); CU(dev_alloc(&t->pieces, 2));
  {
    // isDockerFile scratch (DevTree): sized for the worst case, every file a candidate
    uint64_t scap = (t->shard_bytes + kDfStripe - 1) / kDfStripe + ns + 1;  // stripes over every file body, rounded up per file
    if (scap > 0xFFFFFFF0ull) return fail(C, M2K_ERR_LIMIT, "shard too large for the Dockerfile stripe table");
    t->stripe_cap = (uint32_t)scap;
    const uint64_t cap = scap / kDfTileStripes + 2;
    t->tile_cap = (uint32_t)cap;
    CU(dev_alloc(&t->head_list, ns)); CU(dev_alloc(&t->cand_file, ns)); CU(dev_alloc(&t->cand_stripe0, ns)); CU(dev_alloc(&t->cand_done, ns));
    CU(dev_alloc(&t->cand_err, ns)); CU(dev_alloc(&t->cand_body, ns)); CU(dev_alloc(&t->cand_esc, ns));
    CU(arena_alloc(C, t, 1, (void**)&t->stripe_state, scap)); CU(dev_alloc(&t->tile_state, cap));
    CU(cudaMemsetAsync(t->tile_state, 0, 4ull * cap, C->stream));  // once: pass numbers start at 1
  }
  {
    // the slab (DevTree::slab): zero region, the list part of `special`, all-ones region
    auto up = [](size_t x) { return align256(x); };
    const size_t nd4 = up(4ull * (nd ? nd : 1));
    size_t off = 0;
    const size_t o_direct = off; off += nd4;
    const size_t o_content = off; off += nd4;
    const size_t o_subtree = off; off += nd4;
    const size_t o_flags = off; off += up(nf ? nf : 1);
    const size_t o_ctr = off; off += up(sizeof(DfCounters));
    const size_t o_special = off; off += up(4ull * ((size_t)kSpecialCap + 1));
    t->slab_zero_bytes = o_special + 4;  // up to and including special[0], the count
    t->slab_ff_off = off;
    const size_t o_war = off; off += nd4;
    const size_t o_main = off; off += nd4;
    const size_t o_main_sub = off; off += nd4;
    t->slab_ff_bytes = off - t->slab_ff_off;
    const size_t o_mailbox = off; off += up(sizeof(Mailbox));
    const size_t o_all_pieces = off; off += up(2 * kMaxRanks * sizeof(PieceSummary));
    t->slab_bytes = off;
    CU(cudaMalloc((void**)&t->slab, off));
    t->mailbox = (Mailbox*)(t->slab + o_mailbox); t->all_pieces = (PieceSummary*)(t->slab + o_all_pieces);
    CU(cudaMemsetAsync(t->mailbox, 0, sizeof(Mailbox), C->stream));
    CU(cudaMemsetAsync(t->all_pieces, 0xFF, 2 * kMaxRanks * sizeof(PieceSummary), C->stream));
    t->dir_direct = (uint32_t*)(t->slab + o_direct); t->dir_content = (uint32_t*)(t->slab + o_content);
    t->dir_subtree = (uint32_t*)(t->slab + o_subtree); t->file_flags = t->slab + o_flags;
    t->df_counters = (DfCounters*)(t->slab + o_ctr); t->special = (uint32_t*)(t->slab + o_special);
    t->dir_first_war = (uint32_t*)(t->slab + o_war); t->dir_main = (uint32_t*)(t->slab + o_main);
    t->dir_main_sub = (uint32_t*)(t->slab + o_main_sub);
  }
  {
    // what the walk produces early, in the order of the host result (one copy down): options, states, ignore flags
    const size_t nd4 = align256(4ull * (nd ? nd : 1)), nd1 = align256(nd ? nd : 1);
    CU(cudaMalloc((void**)&t->early, nd4 + 2 * nd1));
    t->dir_opts = (uint32_t*)t->early; t->dir_state = t->early + nd4; t->dir_ignore = t->early + nd4 + nd1;
    t->early_bytes = nd4 + 2 * nd1;
  }
  CU(dev_alloc(&t->dir_descend, nd));

  dt->ctx = C;
  dt->stage.pool = C;
  dt->h_entry_feat = (uint32_t*)pinned_take(C, 4ull * (ne ? ne : 1));
  dt->h_special = (uint32_t*)pinned_take(C, 4ull * (kSpecialCap + 1));
  dt->h_results = (uint8_t*)pinned_take(C, 5 * align256(4ull * (nd ? nd : 1)) + 2 * align256(nd ? nd : 1) + nf + 256);
  if (!dt->h_entry_feat || !dt->h_special || !dt->h_results) return fail(C, M2K_ERR_NOMEM, "out of pinned host memory for the result buffers");
  t->special_cap = kSpecialCap;
  std::vector<uint32_t> chunk_file(t->n_chunks ? t->n_chunks : 1);
  for (uint32_t i = 0; i < ns; i++) std::fill(chunk_file.begin() + chunk_first[i], chunk_file.begin() + chunk_first[i + 1], i);
  CU(cudaMemcpyAsync(t->chunk_first, chunk_first.data(), 4ull * ((size_t)ns + 1), cudaMemcpyHostToDevice, C->stream));
  CU(cudaMemcpyAsync(t->chunk_file, chunk_file.data(), 4ull * t->n_chunks, cudaMemcpyHostToDevice, C->stream));
  CU(cudaStreamSynchronize(C->stream));  // chunk_first / chunk_file are stack-owned host vectors
  if (getenv("M2K_DEBUG_TIMING")) fprintf(stderr, "[m2k] upload: set-up (layout, allocations, pinned result buffers) %.2f ms\n", now_ms() - t_setup);
  if (int rc = copy_shard(C, dt, tree)) return rc;
  guard.keep = true;
  *out = dt;
  return M2K_OK;
}

// metadata and the shard's bytes, host to device, into the buffers of an existing device tree; timed
static int copy_shard(Ctx* C, m2k_devtree* dt, const m2k_tree* tree) {
  DevTree* t = &dt->t;
  const uint32_t nd = t->n_dirs, ne = t->n_entries, nf = t->n_files;
  cudaStream_t s = C->stream;
  CU(cudaEventRecord(C->ev[0], s));
  auto H2D = [&](void* d, const void* h, size_t bytes) { return bytes ? cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s) : cudaSuccess; };
  CU(H2D(t->dir_parent, tree->dir_parent, 4ull * nd)); CU(H2D(t->dir_entry, tree->dir_entry, 4ull * nd));
  CU(H2D(t->dir_depth, tree->dir_depth, 4ull * nd)); CU(H2D(t->dir_flags, tree->dir_flags, nd));
  CU(H2D(t->entry_dir, tree->entry_dir, 4ull * ne)); CU(H2D(t->entry_name_off, tree->entry_name_off, 4ull * ((size_t)ne + 1)));
  CU(H2D(t->entry_child, tree->entry_child, 4ull * ne)); CU(H2D(t->entry_type, tree->entry_type, ne));
  CU(H2D(t->names, tree->names, tree->n_name_bytes));
  CU(H2D(t->file_entry, tree->file_entry, 4ull * nf)); CU(H2D(t->file_off, tree->file_off, 8ull * nf));
  CU(H2D(t->file_len, tree->file_len, 8ull * nf));
  if (tree->bytes && t->shard_bytes > 16) { if (int rc = copy_to_device(C, s, t->bytes, tree->bytes + t->shard_byte0, t->shard_bytes - 16)) return rc; }
  CU(cudaMemsetAsync(t->bytes + (t->shard_bytes - 16), 0, 16, s));
  CU(cudaEventRecord(C->ev[1], s));
  CU(cudaStreamSynchronize(s));
  CU(cudaEventElapsedTime(&dt->ms_h2d, C->ev[0], C->ev[1]));
  return M2K_OK;
}

// The same shard again from the host (a tree whose files changed in place, or a plugin that keeps its
// device buffers between calls): no allocation, only the copies of m2k_upload.
extern "C" int m2k_reupload(m2k_ctx* ctx, m2k_devtree* dt, const m2k_tree* tree) {
  if (!ctx || !dt || !tree) return fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_reupload: NULL argument");
  Ctx* C = &ctx->c;
  DevTree* t = &dt->t;
  if (tree->n_dirs != t->n_dirs || tree->n_entries != t->n_entries || tree->n_files != t->n_files)
    return fail(C, M2K_ERR_ARG, "m2k_reupload: the tree has another shape than the one this device tree was made for");
  CU(cudaSetDevice(C->device));
  CU(cudaStreamSynchronize(C->stream));
  t->host = tree;
  return copy_shard(C, dt, tree);
}

extern "C" int m2k_upload(m2k_ctx* ctx, const m2k_tree* tree, uint32_t file_begin, uint32_t file_end, m2k_devtree** out) {
  if (!ctx) return fail(nullptr, M2K_ERR_ARG, "ctx is NULL");
  if (!tree || !out || file_begin > file_end || file_end > tree->n_files) return fail(&ctx->c, M2K_ERR_ARG, "m2k_upload: bad argument");
  return upload_impl(ctx, tree, file_begin, file_end, 0, M2K_NONE, out);
}

extern "C" uint64_t m2k_chunk_count(const m2k_tree* tree) {
  uint64_t n = 0;
  if (tree) for (uint32_t f = 0; f < tree->n_files; f++) n += file_chunks(tree, f);
  return n;
}

extern "C" int m2k_upload_chunks(m2k_ctx* ctx, const m2k_tree* tree, uint64_t chunk_begin, uint64_t chunk_end, m2k_devtree** out) {
  if (!ctx) return fail(nullptr, M2K_ERR_ARG, "ctx is NULL");
  if (!tree || !out || chunk_begin > chunk_end) return fail(&ctx->c, M2K_ERR_ARG, "m2k_upload_chunks: bad argument");
  // locate the two ends in the global chunk numbering (files in id order, ceil(len / 64 KiB) chunks each)
  uint32_t f0 = 0, fl = 0, first_skip = 0, last_end = M2K_NONE;
  bool have0 = false, have1 = false;
  uint64_t at = 0;
  for (uint32_t f = 0; f < tree->n_files && !have1; f++) {
    const uint64_t nc = file_chunks(tree, f);
    if (nc == 0) continue;
    if (!have0 && chunk_begin < at + nc) { have0 = true; f0 = f; first_skip = (uint32_t)(chunk_begin - at); }
    if (have0 && chunk_end <= at + nc) { have1 = true; fl = f; last_end = chunk_end - at == nc ? M2K_NONE : (uint32_t)(chunk_end - at); }
    at += nc;
  }
  if (chunk_begin == chunk_end) return upload_impl(ctx, tree, 0, 0, 0, M2K_NONE, out);
  if (!have0 || !have1) return fail(&ctx->c, M2K_ERR_ARG, "m2k_upload_chunks: chunk range [%llu, %llu) is outside the tree's %llu chunks",
                                    (unsigned long long)chunk_begin, (unsigned long long)chunk_end, (unsigned long long)m2k_chunk_count(tree));
  return upload_impl(ctx, tree, f0, fl + 1, first_skip, last_end, out);
}

// What a chunk costs the byte scan: its bytes, plus a fixed part (a warp's set-up, the file lookup,
// the record) worth about 4 KiB of streaming -- measured on config D, where a million files of a
// few hundred bytes take as long as ~5 GB of large ones.  Balancing bytes alone would hand one GPU
// all the tiny files; balancing chunks alone would hand another all the multi-GiB ones.
static constexpr uint64_t kChunkFixedCost = 4096;
static inline uint64_t file_cost(const m2k_tree* tree, uint32_t f) { return tree->file_len[f] + kChunkFixedCost * file_chunks(tree, f); }

extern "C" uint64_t m2k_root_extra_bytes(const m2k_tree* tree) {
  // What the rank that assembles the result does besides scanning its shard, expressed as bytes of scanning
  // (7 TB/s).  The other ranks stay one pass ahead of their hosts (m2k_exchange_wait_previous), so a pass takes
  // max(root's scan + its tail, a rank's scan + its per-file kernels); the root's tail is the longer one by its
  // stitching, k_content_up, result copies and host work, and its scan starts behind the name kernel and shares
  // the SMs with its directory-level kernels.  Swept on 2 x B200 (0.1 / 0.25 / 0.4 ms: config E 7.42 / 7.46 /
  // 7.56 ms, D 4.12 / 4.16 / 4.28, 2 x C 14.50 / 14.41 / 14.41): 0.1 ms plus 0.06 ms per million entries.
  if (!tree) return 0;
  return (uint64_t)(7.0e12 /* bytes per second */ * (0.10e-3 + 0.06e-9 * (double)tree->n_entries) /* seconds */);
}

extern "C" int m2k_plan_shards(const m2k_tree* tree, uint32_t n_ranks, uint64_t split_min_bytes, uint64_t root_extra_bytes,
                               uint64_t* chunk_cuts) {
  if (!tree || !n_ranks || !chunk_cuts) return fail(nullptr, M2K_ERR_ARG, "m2k_plan_shards: bad argument");
  uint64_t total_cost = 0, total_chunks = 0;
  for (uint32_t f = 0; f < tree->n_files; f++) { total_cost += file_cost(tree, f); total_chunks += file_chunks(tree, f); }
  // the discount is meant for shards that take milliseconds to scan: never more than half of an even share
  root_extra_bytes = std::min<uint64_t>(root_extra_bytes, total_cost / n_ranks / 2);
  chunk_cuts[0] = 0;
  uint32_t f = 0;
  uint64_t at_cost = 0, at_chunk = 0;  // cost / chunks of the files before f
  for (uint32_t r = 1; r < n_ranks; r++) {
    // rank 0 is charged root_extra_bytes on top of its shard, so every cut moves left by that much
    const unsigned __int128 ideal = ((unsigned __int128)total_cost + root_extra_bytes) * r / n_ranks;
    const uint64_t target = ideal > root_extra_bytes ? (uint64_t)(ideal - root_extra_bytes) : 0;
    while (f < tree->n_files && at_cost + file_cost(tree, f) <= target) { at_cost += file_cost(tree, f); at_chunk += file_chunks(tree, f); f++; }
    uint64_t cut = at_chunk;
    if (f < tree->n_files && target > at_cost) {  // the target falls inside file f
      const uint64_t nc = file_chunks(tree, f), into = target - at_cost;
      if (tree->file_len[f] >= split_min_bytes) cut = at_chunk + std::min<uint64_t>(nc, into / (kChunkBytes + kChunkFixedCost));
      else cut = into * 2 < file_cost(tree, f) ? at_chunk : at_chunk + nc;  // small file: whole, to the nearer side
    }
    chunk_cuts[r] = std::max(cut, chunk_cuts[r - 1]);
  }
  chunk_cuts[n_ranks] = total_chunks;
  return M2K_OK;
}

extern "C" int m2k_synth_fill_device(m2k_ctx* ctx, m2k_devtree* dt, uint64_t seed) {
  if (!ctx || !dt) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  DevTree* t = &dt->t;
  const TreeOwner* o = static_cast<const TreeOwner*>(t->host->owner);
  if (!o || !o->synthetic) return fail(C, M2K_ERR_ARG, "m2k_synth_fill_device: not a synthetic tree");
  if (seed != o->synth_seed) return fail(C, M2K_ERR_ARG, "seed differs from the tree's seed");
  CU(cudaSetDevice(C->device));
  if (!t->synth_class) {
    CU(dev_alloc(&t->synth_class, t->n_files)); CU(dev_alloc(&t->synth_inject, t->n_files));
    CU(cudaMemcpyAsync(t->synth_class, o->synth_class.data(), t->n_files, cudaMemcpyHostToDevice, C->stream));
    CU(cudaMemcpyAsync(t->synth_inject, o->synth_inject.data(), 8ull * t->n_files, cudaMemcpyHostToDevice, C->stream));
  }
  launch_synth_fill(C, t, seed);
  CU(cudaGetLastError());
  for (size_t k = 0; k < o->literal_file.size(); k++) {
    uint32_t f = o->literal_file[k];
    if (f < t->file_begin || f >= t->file_end) continue;
    uint64_t off = t->host->file_off[f] - t->shard_byte0;
    uint64_t padded = (t->host->file_len[f] + 15) & ~15ull;
    CU(cudaMemsetAsync(t->bytes + off, 0, padded, C->stream));
    CU(cudaMemcpyAsync(t->bytes + off, o->literal_data[k].data(), o->literal_data[k].size(), cudaMemcpyHostToDevice, C->stream));
  }
  CU(cudaStreamSynchronize(C->stream));
  return M2K_OK;
}

// ---- phase A --------------------------------------------------------------------------------
extern "C" int m2k_scan(m2k_ctx* ctx, m2k_devtree* dt, const m2k_detect_params* params) {
  if (!ctx || !dt) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  DevTree* t = &dt->t;
  CU(cudaSetDevice(C->device));
  cudaStream_t s = C->stream;
  const bool content = !(params && params->skip_content);
  CU(cudaStreamWaitEvent(s, C->ev_feat, 0));   // a previous pass's copy of the name features may still be reading them
  CU(cudaStreamWaitEvent(s, C->ev_early, 0));  // ... and its directory-level kernels the arrays cleared below
  CU(cudaMemsetAsync(t->slab, 0, t->slab_zero_bytes, s));
  CU(cudaMemsetAsync(t->slab + t->slab_ff_off, 0xFF, t->slab_ff_bytes, s));
  if (++t->epoch >= 0x1FFFFFFFu) {  // the tile table is tagged with the pass number: start over
    CU(cudaMemsetAsync(t->tile_state, 0, 4ull * t->tile_cap, s));
    t->epoch = 1;
  }
  // multi-GPU over peer memory: a rank attached to a root only scans its shard and reduces into the
  // root's arrays; names, ignore rules and the walk are the root's business
  PeerLink& L = t->link;
  const bool remote = L.attached && !L.is_root;
  if (L.attached) {
    L.pass++;
    if (L.is_root) launch_signal(C, &t->mailbox->ready, L.pass, s);  // my arrays are cleared: contributions for this pass may come
  }
  dt->prep.ready = false;
  CU(cudaEventRecord(C->ev[8], s));
  if (!remote) {
    launch_name_features(C, t);
    // the host needs the name features (which entries are .m2kignore files) while the byte scan runs:
    // they go down on the copy stream, so the byte scan starts right behind the name kernel
    CU(cudaEventRecord(C->ev_names, s));
    CU(cudaStreamWaitEvent(C->copy_stream, C->ev_names, 0));
    CU(cudaMemcpyAsync(dt->h_special, t->special, 4ull * (kSpecialCap + 1), cudaMemcpyDeviceToHost, C->copy_stream));
    CU(cudaMemcpyAsync(dt->h_entry_feat, t->entry_feat, 4ull * t->n_entries, cudaMemcpyDeviceToHost, C->copy_stream));
    CU(cudaEventRecord(C->ev_feat, C->copy_stream));
  }
  CU(cudaEventRecord(C->ev[2], s));
  dt->content = content;
  if (content) {
    launch_scan_bytes(C, t, t->chunk_rec, &t->df_counters->scan_ticket);
    CU(cudaEventRecord(C->ev[3], s));
    if (remote) launch_wait_ge(C, &L.mailbox->ready, 1, L.pass, &L.mailbox->abort, s);  // the root has cleared its arrays for this pass
    launch_file_phase(C, t, t->chunk_rec);
    CU(cudaEventRecord(C->ev[9], s));
    launch_dockerfile(C, t);
    CU(cudaEventRecord(C->ev[10], s));
  } else if (remote) {
    CU(cudaMemsetAsync(t->pieces, 0xFF, 2 * sizeof(PieceSummary), s));
  }
  if (remote) {
    launch_contribute(C, t, s);
    launch_signal(C, &L.mailbox->done[L.rank], L.pass, s);
    cudaEvent_t& ev = C->ev_pass[L.pass & 1u];
    if (!ev) CU(cudaEventCreateWithFlags(&ev, cudaEventBlockingSync | cudaEventDisableTiming));
    CU(cudaEventRecord(ev, s));
  }
  CU(cudaGetLastError());
  // no synchronisation here: m2k_finalize overlaps its host work with these kernels
  t->scanned = true;
  return M2K_OK;
}

static int collect_scan_timings(Ctx* C, m2k_devtree* dt) {
  memset(C->phase_ms, 0, sizeof C->phase_ms);
  CU(cudaEventElapsedTime(&C->phase_ms[0], C->ev[8], C->ev[2]));
  if (dt->content) {
    CU(cudaEventElapsedTime(&C->last_scan_ms, C->ev[2], C->ev[3]));
    C->phase_ms[1] = C->last_scan_ms;
    CU(cudaEventElapsedTime(&C->phase_ms[2], C->ev[3], C->ev[9]));
    CU(cudaEventElapsedTime(&C->phase_ms[3], C->ev[9], C->ev[10]));
  }
  return M2K_OK;
}

extern "C" int m2k_partials(m2k_ctx* ctx, m2k_devtree* dt, uint32_t* d_dst_content, uint32_t* d_dst_main) {
  if (!ctx || !dt || !d_dst_content || !d_dst_main) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  CU(cudaMemcpyAsync(d_dst_content, dt->t.dir_content, 4ull * dt->t.n_dirs, cudaMemcpyDeviceToDevice, C->stream));
  CU(cudaMemcpyAsync(d_dst_main, dt->t.dir_main, 4ull * dt->t.n_dirs, cudaMemcpyDeviceToDevice, C->stream));
  CU(cudaStreamSynchronize(C->stream));
  return M2K_OK;
}

// per-file flags of a sharded run: every rank fills only its own file range (the rest stays zero), so
// a byte-wise sum over ranks is the union
extern "C" int m2k_file_flags_get(m2k_ctx* ctx, m2k_devtree* dt, uint8_t* d_dst) {
  if (!ctx || !dt || !d_dst) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  CU(cudaMemcpyAsync(d_dst, dt->t.file_flags, dt->t.n_files, cudaMemcpyDeviceToDevice, C->stream));
  CU(cudaStreamSynchronize(C->stream));
  return M2K_OK;
}
extern "C" int m2k_file_flags_set(m2k_ctx* ctx, m2k_devtree* dt, const uint8_t* d_src) {
  if (!ctx || !dt || !d_src) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  CU(cudaMemcpyAsync(dt->t.file_flags, d_src, dt->t.n_files, cudaMemcpyDeviceToDevice, C->stream));
  return M2K_OK;
}

// the (at most two) files this shard holds only part of: 2 x 32 bytes, device to device
extern "C" int m2k_pieces(m2k_ctx* ctx, m2k_devtree* dt, void* d_dst) {
  if (!ctx || !dt || !d_dst) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  if (dt->content) CU(cudaMemcpyAsync(d_dst, dt->t.pieces, 2 * sizeof(PieceSummary), cudaMemcpyDeviceToDevice, C->stream));
  else CU(cudaMemsetAsync(d_dst, 0xFF, 2 * sizeof(PieceSummary), C->stream));
  CU(cudaStreamSynchronize(C->stream));
  return M2K_OK;
}

extern "C" int m2k_merge_pieces(m2k_ctx* ctx, m2k_devtree* dt, const void* d_all_pieces, uint32_t n_ranks) {
  if (!ctx || !dt || !d_all_pieces || !n_ranks) return fail(nullptr, M2K_ERR_ARG, "bad argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  launch_merge_pieces(C, &dt->t, static_cast<const PieceSummary*>(d_all_pieces), 2 * n_ranks);
  CU(cudaGetLastError());
  return M2K_OK;
}

extern "C" uint64_t m2k_devtree_content_bytes(const m2k_devtree* dt) { return dt ? dt->t.shard_content : 0; }

extern "C" int m2k_file_flag(m2k_ctx* ctx, m2k_devtree* dt, uint32_t f, uint8_t* out) {
  if (!ctx || !dt || !out || f >= dt->t.n_files) return fail(nullptr, M2K_ERR_ARG, "bad argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  CU(cudaMemcpyAsync(out, dt->t.file_flags + f, 1, cudaMemcpyDeviceToHost, C->stream));
  CU(cudaStreamSynchronize(C->stream));
  return M2K_OK;
}

extern "C" int m2k_settle_file(m2k_ctx* ctx, m2k_devtree* dt, uint32_t f, uint32_t flags) {
  if (!ctx || !dt || f >= dt->t.n_files || (flags & ~0xFu)) return fail(nullptr, M2K_ERR_ARG, "bad argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  launch_settle_file(C, &dt->t, f, flags);
  CU(cudaGetLastError());
  return M2K_OK;
}

extern "C" int m2k_chunk_locate(const m2k_tree* tree, uint64_t chunk, uint32_t* file, uint32_t* chunk_in_file) {
  if (!tree || !file || !chunk_in_file) return fail(nullptr, M2K_ERR_ARG, "bad argument");
  uint64_t at = 0;
  for (uint32_t f = 0; f < tree->n_files; f++) {
    const uint64_t nc = file_chunks(tree, f);
    if (chunk < at + nc) { *file = f; *chunk_in_file = (uint32_t)(chunk - at); return M2K_OK; }
    at += nc;
  }
  return fail(nullptr, M2K_ERR_ARG, "chunk %llu is past the tree's %llu chunks", (unsigned long long)chunk, (unsigned long long)at);
}

// The exchange of a sharded pass without a host synchronisation: everything this rank contributes is
// copied (device to device) into the caller's buffers on the library's stream, and everything gathered
// is merged from the caller's buffers on that stream.  The caller orders its own stream (the one its
// collectives run on) against m2k_stream() with events -- and keeps the buffers alive and untouched
// until the stream has passed the copies / merges (m2k_finalize synchronises it).
extern "C" int m2k_export_async(m2k_ctx* ctx, m2k_devtree* dt, uint32_t* d_content, uint32_t* d_main, void* d_pieces, uint8_t* d_flags) {
  if (!ctx || !dt || !d_content || !d_main || !d_pieces || !d_flags) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  DevTree* t = &dt->t;
  CU(cudaSetDevice(C->device));
  cudaStream_t s = C->stream;
  CU(cudaMemcpyAsync(d_content, t->dir_content, 4ull * t->n_dirs, cudaMemcpyDeviceToDevice, s));
  CU(cudaMemcpyAsync(d_main, t->dir_main, 4ull * t->n_dirs, cudaMemcpyDeviceToDevice, s));
  if (dt->content) CU(cudaMemcpyAsync(d_pieces, t->pieces, 2 * sizeof(PieceSummary), cudaMemcpyDeviceToDevice, s));
  else CU(cudaMemsetAsync(d_pieces, 0xFF, 2 * sizeof(PieceSummary), s));
  CU(cudaMemcpyAsync(d_flags, t->file_flags, t->n_files, cudaMemcpyDeviceToDevice, s));
  return M2K_OK;
}

extern "C" int m2k_import_async(m2k_ctx* ctx, m2k_devtree* dt, const uint32_t* d_all_content, const uint32_t* d_all_main,
                                const void* d_all_pieces, const uint8_t* d_flags, uint32_t n_ranks) {
  if (!ctx || !dt || !d_all_content || !d_all_main || !d_all_pieces || !d_flags || !n_ranks) return fail(nullptr, M2K_ERR_ARG, "bad argument");
  Ctx* C = &ctx->c;
  DevTree* t = &dt->t;
  CU(cudaSetDevice(C->device));
  launch_merge_partials(C, t, d_all_content, d_all_main, n_ranks);
  CU(cudaMemcpyAsync(t->file_flags, d_flags, t->n_files, cudaMemcpyDeviceToDevice, C->stream));
  launch_merge_pieces(C, t, static_cast<const PieceSummary*>(d_all_pieces), 2 * n_ranks);
  CU(cudaGetLastError());
  return M2K_OK;
}

extern "C" int m2k_merge_partials(m2k_ctx* ctx, m2k_devtree* dt, const uint32_t* d_all_content, const uint32_t* d_all_main, uint32_t n_ranks) {
  if (!ctx || !dt || !d_all_content || !d_all_main || !n_ranks) return fail(nullptr, M2K_ERR_ARG, "bad argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  launch_merge_partials(C, &dt->t, d_all_content, d_all_main, n_ranks);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(C->stream));
  return M2K_OK;
}

// ---- multi-GPU over peer memory ---------------------------------------------------------------
namespace {
struct XHandle {  // what m2k_exchange_open hands to the other ranks (fits m2k_exchange_handle)
  uint32_t magic, n_ranks;
  int32_t device, pid;
  uint32_t n_dirs, n_files;
  uint64_t slab_bytes, off_content, off_main, off_flags, off_mailbox, off_pieces;
  cudaIpcMemHandle_t ipc;
};
static_assert(sizeof(XHandle) <= sizeof(m2k_exchange_handle), "handle blob too small");
constexpr uint32_t kXMagic = 0x4D324B58u;  // "M2KX"

void link_from_base(PeerLink& L, uint8_t* base, const XHandle& h) {
  L.dir_content = (uint32_t*)(base + h.off_content); L.dir_main = (uint32_t*)(base + h.off_main);
  L.file_flags = base + h.off_flags; L.mailbox = (Mailbox*)(base + h.off_mailbox); L.pieces = (PieceSummary*)(base + h.off_pieces);
}
XHandle describe_root(const Ctx* C, const DevTree* t, uint32_t n_ranks) {
  XHandle h;
  memset(&h, 0, sizeof h);
  h.magic = kXMagic; h.n_ranks = n_ranks; h.device = C->device; h.pid = (int32_t)getpid();
  h.n_dirs = t->n_dirs; h.n_files = t->n_files; h.slab_bytes = t->slab_bytes;
  h.off_content = (uint8_t*)t->dir_content - t->slab; h.off_main = (uint8_t*)t->dir_main - t->slab;
  h.off_flags = t->file_flags - t->slab; h.off_mailbox = (uint8_t*)t->mailbox - t->slab; h.off_pieces = (uint8_t*)t->all_pieces - t->slab;
  return h;
}
}  // namespace

extern "C" int m2k_exchange_open(m2k_ctx* ctx, m2k_devtree* dt, uint32_t n_ranks, m2k_exchange_handle* out) {
  if (!ctx || !dt || !out || n_ranks < 1 || n_ranks > kMaxRanks) return fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_exchange_open: bad argument (1..%u ranks)", kMaxRanks);
  Ctx* C = &ctx->c;
  DevTree* t = &dt->t;
  CU(cudaSetDevice(C->device));
  if (int pe = preload_kernels()) return fail(C, M2K_ERR_CUDA, "loading the kernels: %s", cudaGetErrorString((cudaError_t)pe));
  XHandle h = describe_root(C, t, n_ranks);
  cudaError_t e = cudaIpcGetMemHandle(&h.ipc, t->slab);
  if (e != cudaSuccess) { cudaGetLastError(); memset(&h.ipc, 0, sizeof h.ipc); }  // same-process ranks (m2k_exchange_attach_local) do not need it
  memset(out, 0, sizeof *out);
  memcpy(out, &h, sizeof h);
  PeerLink& L = t->link;
  L = PeerLink();
  L.attached = true; L.is_root = true; L.rank = 0; L.n_ranks = n_ranks;
  link_from_base(L, t->slab, h);
  CU(cudaMemsetAsync(t->mailbox, 0, sizeof(Mailbox), C->stream));
  CU(cudaStreamSynchronize(C->stream));
  return M2K_OK;
}

extern "C" int m2k_exchange_attach(m2k_ctx* ctx, m2k_devtree* dt, uint32_t rank, const m2k_exchange_handle* handle) {
  if (!ctx || !dt || !handle) return fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_exchange_attach: NULL argument");
  Ctx* C = &ctx->c;
  DevTree* t = &dt->t;
  XHandle h;
  memcpy(&h, handle, sizeof h);
  if (h.magic != kXMagic || rank == 0 || rank >= h.n_ranks) return fail(C, M2K_ERR_ARG, "m2k_exchange_attach: not a handle of m2k_exchange_open, or rank %u outside 1..%u", rank, h.n_ranks - 1);
  if (h.n_dirs != t->n_dirs || h.n_files != t->n_files) return fail(C, M2K_ERR_ARG, "m2k_exchange_attach: the root holds a different tree (%u dirs / %u files, here %u / %u)", h.n_dirs, h.n_files, t->n_dirs, t->n_files);
  if (h.pid == (int32_t)getpid()) return fail(C, M2K_ERR_ARG, "m2k_exchange_attach: the root lives in this process: use m2k_exchange_attach_local");
  CU(cudaSetDevice(C->device));
  if (int pe = preload_kernels()) return fail(C, M2K_ERR_CUDA, "loading the kernels: %s", cudaGetErrorString((cudaError_t)pe));
  void* base = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&base, h.ipc, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(C, M2K_ERR_CUDA, "cudaIpcOpenMemHandle (root on device %d): %s", h.device, cudaGetErrorString(e)); }
  PeerLink& L = t->link;
  if (L.ipc_base) cudaIpcCloseMemHandle(L.ipc_base);
  L = PeerLink();
  L.attached = true; L.is_root = false; L.rank = rank; L.n_ranks = h.n_ranks; L.ipc_base = base;
  link_from_base(L, static_cast<uint8_t*>(base), h);
  return M2K_OK;
}

extern "C" int m2k_exchange_attach_local(m2k_ctx* ctx, m2k_devtree* dt, uint32_t rank, m2k_ctx* root_ctx, m2k_devtree* root_dt) {
  if (!ctx || !dt || !root_ctx || !root_dt) return fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_exchange_attach_local: NULL argument");
  Ctx* C = &ctx->c;
  DevTree *t = &dt->t, *rt = &root_dt->t;
  if (!rt->link.attached || !rt->link.is_root) return fail(C, M2K_ERR_ARG, "m2k_exchange_attach_local: m2k_exchange_open was not called on the root");
  if (rank == 0 || rank >= rt->link.n_ranks) return fail(C, M2K_ERR_ARG, "m2k_exchange_attach_local: rank %u outside 1..%u", rank, rt->link.n_ranks - 1);
  if (rt->n_dirs != t->n_dirs || rt->n_files != t->n_files) return fail(C, M2K_ERR_ARG, "m2k_exchange_attach_local: the root holds a different tree");
  CU(cudaSetDevice(C->device));
  if (int pe = preload_kernels()) return fail(C, M2K_ERR_CUDA, "loading the kernels: %s", cudaGetErrorString((cudaError_t)pe));
  if (C->device != root_ctx->c.device) {
    int can = 0;
    CU(cudaDeviceCanAccessPeer(&can, C->device, root_ctx->c.device));
    if (!can) return fail(C, M2K_ERR_CUDA, "device %d cannot access device %d's memory (no NVLink / PCIe peer path)", C->device, root_ctx->c.device);
    cudaError_t e = cudaDeviceEnablePeerAccess(root_ctx->c.device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(C, M2K_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d): %s", root_ctx->c.device, cudaGetErrorString(e));
    cudaGetLastError();
  }
  PeerLink& L = t->link;
  L = PeerLink();
  L.attached = true; L.is_root = false; L.rank = rank; L.n_ranks = rt->link.n_ranks;
  link_from_base(L, rt->slab, describe_root(&root_ctx->c, rt, rt->link.n_ranks));
  return M2K_OK;
}

// the root, after its own m2k_scan: take in what the other ranks delivered for this pass
extern "C" int m2k_exchange_collect(m2k_ctx* ctx, m2k_devtree* dt) {
  if (!ctx || !dt) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  DevTree* t = &dt->t;
  PeerLink& L = t->link;
  if (!L.attached || !L.is_root) return fail(C, M2K_ERR_ARG, "m2k_exchange_collect: not the root of an exchange");
  CU(cudaSetDevice(C->device));
  cudaStream_t s = C->stream;
  if (dt->content) CU(cudaMemcpyAsync(t->all_pieces, t->pieces, 2 * sizeof(PieceSummary), cudaMemcpyDeviceToDevice, s));
  else CU(cudaMemsetAsync(t->all_pieces, 0xFF, 2 * sizeof(PieceSummary), s));
  launch_wait_ge(C, &t->mailbox->done[1], L.n_ranks - 1, L.pass, &t->mailbox->abort, s);
  launch_merge_pieces(C, t, t->all_pieces, 2 * L.n_ranks);
  CU(cudaGetLastError());
  return M2K_OK;
}

// a rank that only scans: wait until its part of the pass has been delivered (blocks, does not spin)
extern "C" int m2k_exchange_wait(m2k_ctx* ctx, m2k_devtree* dt) {
  if (!ctx || !dt) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  CU(cudaSetDevice(C->device));
  if (!C->ev_idle) CU(cudaEventCreateWithFlags(&C->ev_idle, cudaEventBlockingSync | cudaEventDisableTiming));
  CU(cudaEventRecord(C->ev_idle, C->stream));
  CU(cudaEventSynchronize(C->ev_idle));
  if (int rc = collect_scan_timings(C, dt)) return rc;
  return M2K_OK;
}

// An attached rank may queue its next pass before the previous one has been delivered: everything it does is
// ordered on its own stream, and what it writes into the root's memory waits (on the device) for the root's
// "cleared for this pass" flag.  Its host then only has to stay one pass behind: this waits for the pass BEFORE
// the last m2k_scan, so the rank's next byte scan starts the moment its GPU is free instead of after a host
// wake-up (measured: 0.2-0.4 ms per pass at 2 and 8 GPUs).
extern "C" int m2k_exchange_wait_previous(m2k_ctx* ctx, m2k_devtree* dt) {
  if (!ctx || !dt) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  const PeerLink& L = dt->t.link;
  if (!L.attached || L.is_root || L.pass < 2) return M2K_OK;
  CU(cudaSetDevice(C->device));
  cudaEvent_t ev = C->ev_pass[(L.pass - 1) & 1u];
  if (ev) CU(cudaEventSynchronize(ev));
  return M2K_OK;
}

// ---- phase C --------------------------------------------------------------------------------
namespace {

const uint8_t* host_file_bytes(const m2k_tree* t, uint32_t f, uint64_t* len) {
  *len = t->file_len[f];
  // synthetic trees keep their rule files as literals (their bytes may live on another rank's shard);
  // the generator numbers files in increasing order, so literal_file is sorted
  const TreeOwner* o = static_cast<const TreeOwner*>(t->owner);
  if (o && !o->literal_file.empty()) {
    auto it = std::lower_bound(o->literal_file.begin(), o->literal_file.end(), f);
    if (it != o->literal_file.end() && *it == f) return reinterpret_cast<const uint8_t*>(o->literal_data[it - o->literal_file.begin()].data());
  }
  if (t->bytes) return t->bytes + t->file_off[f];
  return nullptr;
}

// one block: the pointer array followed by the characters.  The block belongs to the device tree and
// is reused from pass to pass (grow-only): handing a fresh 600 KB block to every result of a 10 000-rule
// tree cost more in mmap / munmap than the whole host half of the pass.
char** list_from_arenas(const std::vector<PathArena>& arenas, uint8_t flag, uint32_t* count, ListBuf* buf) {
  size_t n = 0, chars = 0;
  for (const PathArena& A : arenas)
    for (size_t i = 0; i < A.size(); i++)
      if (A.flag[i] == flag) { n++; chars += A.off[i + 1] - A.off[i] + 1; }
  size_t head = (n ? n : 1) * sizeof(char*);
  if (head + chars + 1 > buf->cap) {
    free(buf->p);
    buf->cap = (head + chars + 1) * 2;
    buf->p = static_cast<char*>(malloc(buf->cap));
  }
  char** out = reinterpret_cast<char**>(buf->p);
  char* w = reinterpret_cast<char*>(out) + head;
  out[0] = nullptr;
  size_t k = 0;
  for (const PathArena& A : arenas)
    for (size_t i = 0; i < A.size(); i++)
      if (A.flag[i] == flag) {
        size_t len = A.off[i + 1] - A.off[i];
        out[k++] = w;
        memcpy(w, A.chars.data() + A.off[i], len);
        w[len] = 0;
        w += len + 1;
      }
  *count = (uint32_t)n;
  return out;
}

}  // namespace

extern "C" void m2k_result_free(m2k_result* r) {
  if (!r) return;
  // the per-directory / per-file / per-entry arrays live in pinned buffers of the device tree (the
  // D2H copies land there directly); a result of m2k_detect owns what is left of its device tree
  // (the two ignore lists live in blocks of the device tree as well)
  if (r->owner) {
    m2k_devtree* dt = static_cast<m2k_devtree*>(r->owner);
    if (dt->pending == r) dt->pending = nullptr;
    free_devtree_shell(dt);
  }
  free(r);
}

// getIgnorePaths (any2kube.go:151-181) + the lookup structure of IsStringPresent (utils.go:308-315):
// host work that needs only the name features, so it runs while the byte scan is still on the GPU.
static unsigned host_threads(size_t work_items, size_t per_thread) {
  unsigned nt = std::thread::hardware_concurrency();
  if (nt == 0) nt = 4;
  if (nt > 32) nt = 32;
  if (nt > work_items / per_thread + 1) nt = (unsigned)(work_items / per_thread + 1);
  return nt;
}

// A small persistent team of host threads for the parse / fold / insert phases of m2k_prepare:
// spawning up to 32 threads four times per pass cost more than the work itself.  Created on first
// use, kept for the life of the process (never joined: the workers sleep on a condition variable).
class HostTeam {
 public:
  static HostTeam& get() { static HostTeam* t = new HostTeam(); return *t; }
  void run(unsigned nt, const std::function<void(unsigned)>& fn) {
    if (nt <= 1) { fn(0); return; }
    if (nt > workers_ + 1) nt = workers_ + 1;
    // one job at a time: contexts driven from different host threads share this team
    std::lock_guard<std::mutex> one(run_m_);
    {
      std::lock_guard<std::mutex> lk(m_);
      job_ = &fn; active_ = nt - 1; remaining_ = nt - 1; gen_++;
    }
    cv_.notify_all();
    fn(0);
    std::unique_lock<std::mutex> lk(m_);
    done_.wait(lk, [&] { return remaining_ == 0; });
    job_ = nullptr;
  }
  unsigned size() const { return workers_ + 1; }

 private:
  HostTeam() {
    unsigned hw = std::thread::hardware_concurrency();
    if (hw == 0) hw = 4;
    workers_ = std::min(hw, 32u) - 1;
    for (unsigned i = 1; i <= workers_; i++) std::thread([this, i] { loop(i); }).detach();
  }
  void loop(unsigned id) {
    uint64_t seen = 0;
    for (;;) {
      const std::function<void(unsigned)>* job = nullptr;
      {
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { return gen_ != seen; });
        seen = gen_;
        if (id <= active_) job = job_;
      }
      if (!job) continue;
      (*job)(id);
      std::lock_guard<std::mutex> lk(m_);
      if (--remaining_ == 0) done_.notify_one();
    }
  }
  std::mutex m_, run_m_;
  std::condition_variable cv_, done_;
  const std::function<void(unsigned)>* job_ = nullptr;
  unsigned workers_ = 0, active_ = 0, remaining_ = 0;
  uint64_t gen_ = 0;
};

template <typename F>
static void run_threads(unsigned nt, F&& fn) {
  std::function<void(unsigned)> f = fn;
  HostTeam::get().run(nt, f);
}

// The host half of m2k_prepare: P.ign_entries (the .m2kignore entries, any order) -> the two cleaned
// lists (P.arenas) and the staged blob (folded runes, depth-1 prefix, hash table) in S.  No CUDA
// calls except the staging allocation.  Returns 0, or an M2K_ERR_* with *err set.
static void parallel_memcpy(void* dst, const void* src, size_t n) {
  const unsigned nt = std::min<unsigned>(host_threads(n, 1u << 20), 16);
  run_threads(nt, [&](unsigned tid) {
    size_t lo = n * tid / nt, hi = n * (tid + 1) / nt;
    memcpy(static_cast<char*>(dst) + lo, static_cast<const char*>(src) + lo, hi - lo);
  });
}

static int prepare_host(const m2k_tree* tree, const m2k_detect_params* params, IgnorePrep& P, Staging& S, std::string* err) {
  double h1 = now_ms();
  const std::string root = tree->root_path ? tree->root_path : "";
  const std::string clean_root = path_clean(root);
  std::vector<uint32_t>& ign_entries = P.ign_entries;
  sort_walk_order(tree, &ign_entries);  // getIgnorePaths reads the files in the order GetFilesByName found them
  const size_t nfiles = ign_entries.size();
  double h2 = now_ms();
  // parse the files in parallel: thread k takes a contiguous run of files (in Walk order) into its own
  // arena, so the arenas in thread order are the reference's two lists in order
  const unsigned nt = host_threads(nfiles, 64);
  P.arenas.resize(nt + 1);  // the last arena holds the preContainerized paths
  for (PathArena& A : P.arenas) { A.chars.clear(); A.off.clear(); A.flag.clear(); A.missing = false; }
  run_threads(nt, [&](unsigned tid) {
    PathArena& A = P.arenas[tid];
    A.off.push_back(0);
    size_t lo = nfiles * tid / nt, hi = nfiles * (tid + 1) / nt;
    for (size_t k = lo; k < hi; k++) {
      uint32_t e = ign_entries[k];
      uint32_t f = tree->entry_child[e];
      if ((tree->entry_type[e] >> 4) != M2K_T_REG || f == M2K_NONE) continue;  // os.Open / read fails: no lines
      uint64_t len;
      const uint8_t* data = host_file_bytes(tree, f, &len);
      if (!data) { if (len) A.missing = true; continue; }
      uint32_t d = tree->entry_dir[e];
      parse_ignore_file_arena(d == 0 ? clean_root : fast_dir_path(tree, d, clean_root), data, len, &A);
    }
  });
  {
    PathArena& A = P.arenas[nt];
    A.off.push_back(0);
    for (uint32_t i = 0; params && i < params->n_pre; i++) {
      const char* pp = params->pre_paths[i] ? params->pre_paths[i] : "";
      A.chars.insert(A.chars.end(), pp, pp + strlen(pp));
      A.off.push_back((uint32_t)A.chars.size());
      A.flag.push_back(4);
    }
  }
  size_t total = 0, total_chars = 0;
  std::vector<size_t> abase(P.arenas.size() + 1, 0), cbase(P.arenas.size() + 1, 0);
  for (size_t k = 0; k < P.arenas.size(); k++) {
    if (P.arenas[k].missing) { *err = ".m2kignore content is not available on the host (tree->bytes is NULL)"; return M2K_ERR_ARG; }
    total += P.arenas[k].size();
    total_chars += P.arenas[k].chars.size();
    abase[k + 1] = total;
    cbase[k + 1] = total_chars;
  }
  double h3 = now_ms();
  // exact hash table over the case-folded entries; flags: 1 ignoreDirectories, 2 ignoreContents,
  // 4 preContainerized.  Fold + hash in parallel, straight into the staging buffer: an entry's runes go
  // where its characters are in the concatenated arenas (a rune is at least one byte), so no thread
  // needs another's count.  Then insert in parallel into 16 regions selected by the top bits of the
  // key (each region is its own linear-probing table, so threads never meet).
  const uint64_t seed0 = 0xcbf29ce484222325ull;
  std::vector<uint32_t> root_runes;
  fold_runes(root, &root_runes);
  // folded prefix shared by all depth-1 directories: Clean(root) + "/" (Join semantics)
  std::string prefix1 = clean_root;
  if (prefix1 == ".") prefix1 = ""; else if (prefix1 != "/") prefix1 += "/";
  std::vector<uint32_t> prefix_runes;
  fold_runes(prefix1, &prefix_runes);
  P.depth1_state = seed0;
  for (uint32_t rr : prefix_runes) P.depth1_state = hash_step(P.depth1_state, rr);
  P.prefix_len = (uint32_t)prefix_runes.size();
  P.root_ignore = 0;
  double h4 = h3, h5 = h3;
  if (total) {
    if (total_chars >= (1ull << 29)) { *err = "the ignore rules hold more than 2^29 characters"; return M2K_ERR_LIMIT; }
    const size_t rb = 4 * total_chars, pb = 4 * prefix_runes.size();
    P.off_prefix = rb;
    P.off_table = (rb + pb + 15) & ~(size_t)15;
    // first guess of the table size (regions evenly filled); corrected once the keys are known
    uint32_t R = 4;
    while ((uint64_t)R * 3 * kIgnoreRegions < (uint64_t)total * 5) R <<= 1;
    if (!S.reserve(P.off_table + sizeof(IgnoreSlot) * R * kIgnoreRegions, 0)) { *err = "out of pinned host memory for the ignore table"; return M2K_ERR_NOMEM; }
    P.folded.resize(total);
    std::vector<Folded>& folded = P.folded;
    uint32_t* runes = static_cast<uint32_t*>(S.p);
    const unsigned nt2 = host_threads(total, 1024);
    std::atomic<uint32_t> root_flags{0};
    run_threads(nt2, [&](unsigned tid) {
      size_t lo = total * tid / nt2, hi = total * (tid + 1) / nt2;
      size_t k = std::upper_bound(abase.begin(), abase.end(), lo) - abase.begin() - 1;
      for (size_t g = lo; g < hi; g++) {
        while (g >= abase[k + 1]) k++;
        const PathArena& A = P.arenas[k];
        const size_t i = g - abase[k];
        const uint8_t* p = reinterpret_cast<const uint8_t*>(A.chars.data()) + A.off[i];
        const size_t n = A.off[i + 1] - A.off[i];
        const uint32_t off = (uint32_t)(cbase[k] + A.off[i]);
        uint32_t* out = runes + off;
        uint32_t len = 0;
        uint64_t h = seed0;
        for (size_t j = 0; j < n;) {
          uint32_t r;
          if (p[j] < 0x80) { r = p[j]; if (r - 'A' < 26u) r += 32; j++; }  // same as fold_rune's ASCII branch
          else { uint32_t w; r = fold_rune(decode_rune(p + j, n - j, &w), m2k_casefold_from, m2k_casefold_to, M2K_CASEFOLD_N); j += w; }
          out[len++] = r;
          h = hash_step(h, r);
        }
        folded[g] = {hash_finish(h), off, len, A.flag[i]};
        if (len == root_runes.size() && std::equal(root_runes.begin(), root_runes.end(), out)) root_flags |= A.flag[i];
      }
    });
    P.root_ignore = (uint8_t)root_flags.load();
    h4 = now_ms();
    uint32_t region_count[kIgnoreRegions] = {0};
    for (size_t g = 0; g < total; g++) region_count[folded[g].key >> (64 - kIgnoreRegionBits)]++;
    uint32_t maxc = 1;
    for (uint32_t c : region_count) maxc = std::max(maxc, c);
    R = 4;
    while ((uint64_t)R * 3 < (uint64_t)maxc * 4) R <<= 1;  // load factor <= 0.75 in the fullest region
    P.table_mask = R - 1;
    const size_t tb = sizeof(IgnoreSlot) * R * kIgnoreRegions;
    P.blob_bytes = P.off_table + tb;
    if (!S.reserve(P.blob_bytes, P.off_table)) { *err = "out of pinned host memory for the ignore table"; return M2K_ERR_NOMEM; }
    runes = static_cast<uint32_t*>(S.p);
    if (pb) memcpy((char*)S.p + P.off_prefix, prefix_runes.data(), pb);
    IgnoreSlot* table = reinterpret_cast<IgnoreSlot*>((char*)S.p + P.off_table);
    const unsigned nt3 = std::min<unsigned>(host_threads(total, 2048), kIgnoreRegions);
    const uint32_t mask = P.table_mask;
    run_threads(nt3, [&](unsigned tid) {
      for (uint32_t region = tid; region < kIgnoreRegions; region += nt3) memset(table + (size_t)region * R, 0, sizeof(IgnoreSlot) * R);
      for (size_t g = 0; g < total; g++) {
        const Folded& fo = folded[g];
        const uint32_t region = (uint32_t)(fo.key >> (64 - kIgnoreRegionBits));
        if (region % nt3 != tid) continue;
        IgnoreSlot* tab = table + (size_t)region * R;
        const uint32_t* rp = runes + fo.off;
        for (uint32_t sidx = (uint32_t)fo.key & mask;; sidx = (sidx + 1) & mask) {
          IgnoreSlot& sl = tab[sidx];
          if (sl.hash == 0) { sl = IgnoreSlot{fo.key, fo.off, slot_pack(fo.len, fo.flag)}; break; }
          if (sl.hash == fo.key && slot_len(sl.len_flags) == fo.len && std::equal(rp, rp + fo.len, runes + sl.off)) {
            sl.len_flags |= slot_pack(0, fo.flag);
            break;
          }
        }
      }
    });
    h5 = now_ms();
  }
  P.ms_part[0] = h2 - h1; P.ms_part[1] = h3 - h2; P.ms_part[2] = h4 - h3; P.ms_part[3] = h5 - h4;
  return M2K_OK;
}

static int prepare_ignore(Ctx* C, m2k_devtree* dt, const m2k_detect_params* params) {
  DevTree* t = &dt->t;
  IgnorePrep& P = dt->prep;
  const m2k_tree* tree = t->host;
  CU(cudaEventSynchronize(C->ev_feat));  // name features and the special-entry list are on the host now
  CU(cudaEventSynchronize(C->ev_blob));  // the staging buffer is free again
  double h0 = now_ms();
  P.reset();
  // the .m2kignore files: from the compacted list of special entries, or a scan of all features if it overflowed
  P.ign_entries.clear();
  const uint32_t n_special = dt->h_special[0];
  if (n_special <= kSpecialCap) {
    for (uint32_t i = 0; i < n_special; i++) {
      uint32_t e = dt->h_special[1 + i];
      if (dt->h_entry_feat[e] & (1u << M2K_N_IGNOREFILE)) P.ign_entries.push_back(e);
    }
  } else {
    for (uint32_t e = 0; e < t->n_entries; e++) if (dt->h_entry_feat[e] & (1u << M2K_N_IGNOREFILE)) P.ign_entries.push_back(e);
  }
  std::string err;
  if (int rc = prepare_host(tree, params, P, dt->stage, &err)) return fail(C, rc, "%s", err.c_str());
  if (P.blob_bytes) {
    if (P.blob_bytes > t->ign_blob_cap) {  // grow-only buffer kept with the device tree
      if (t->ign_blob) CU(cudaFree(t->ign_blob));
      t->ign_blob = nullptr; t->ign_blob_cap = 0;
      CU(cudaMalloc(&t->ign_blob, P.blob_bytes * 2));
      t->ign_blob_cap = P.blob_bytes * 2;
    }
    // on the copy stream: queued on `s` it would only start after the byte scan (and not before an
    // earlier walk, if there was one, has finished reading the table it replaces)
    CU(cudaStreamWaitEvent(C->copy_stream, C->ev_early, 0));
    CU(cudaMemcpyAsync(t->ign_blob, dt->stage.p, P.blob_bytes, cudaMemcpyHostToDevice, C->copy_stream));
    CU(cudaEventRecord(C->ev_blob, C->copy_stream));
    P.d_runes = (const uint32_t*)t->ign_blob;
    P.d_prefix = (const uint32_t*)((char*)t->ign_blob + P.off_prefix);
    P.d_table = (const IgnoreSlot*)((char*)t->ign_blob + P.off_table);
  }
  double h6 = now_ms();
  // the result shell: its arrays are the device tree's pinned buffers
  {
    if (dt->pending) { m2k_result* old = dt->pending; dt->pending = nullptr; m2k_result_free(old); }
    const uint32_t nd = t->n_dirs, ne = t->n_entries, nf = t->n_files;
    m2k_result* r = static_cast<m2k_result*>(calloc(1, sizeof(m2k_result)));
    r->n_dirs = nd; r->n_files = nf; r->n_entries = ne;
    // layout of dt->h_results: the device's early block (options, states, ignore flags) and its
    // [dir_subtree | file_flags] stretch are mirrored, so that each comes down in one copy
    const size_t nd4 = align256(4ull * (nd ? nd : 1)), nd1 = align256(nd ? nd : 1);
    uint8_t* b = dt->h_results;
    r->dir_opts = (uint32_t*)b; r->dir_state = b + nd4; r->dir_ignore = b + nd4 + nd1; b += nd4 + 2 * nd1;
    r->dir_direct = (uint32_t*)b; b += nd4;
    r->dir_main_file = (uint32_t*)b; b += nd4;
    r->service_dirs = (uint32_t*)b; b += nd4;
    r->dir_subtree = (uint32_t*)b; b += nd4;
    r->file_flags = b;
    r->entry_feat = dt->h_entry_feat;
    r->ignore_dirs = list_from_arenas(P.arenas, 1, &r->n_ignore_dirs, &dt->lists[0]);
    r->ignore_contents = list_from_arenas(P.arenas, 2, &r->n_ignore_contents, &dt->lists[1]);
    dt->pending = r;
  }
  // The directory-level phases need the name features and the ignore table, nothing from the byte
  // scan: queue them now, on the high-priority side stream, and send their results down.  What is
  // left for m2k_finalize is the content bits and the main-file ids (k_content_up).
  {
    m2k_result* r = dt->pending;
    const uint32_t nd = t->n_dirs;
    cudaStream_t ss = C->side_stream;
    P.sh_flavor = params ? params->sh_flavor : M2K_SH_BASH;
    CU(cudaStreamWaitEvent(ss, C->ev_names, 0));
    CU(cudaStreamWaitEvent(ss, C->ev_blob, 0));
    CU(cudaEventRecord(C->ev_side0, ss));
    launch_walk(C, t, P.sh_flavor, P.d_table, P.table_mask, P.d_runes, P.d_prefix, P.prefix_len, P.depth1_state, P.root_ignore, ss);
    CU(cudaEventRecord(C->ev_side1, ss));
    CU(cudaMemcpyAsync(r->dir_opts, t->early, t->early_bytes, cudaMemcpyDeviceToHost, ss));  // options, states, ignore flags
    CU(cudaMemcpyAsync(r->dir_direct, t->dir_direct, 4ull * nd, cudaMemcpyDeviceToHost, ss));
    CU(cudaEventRecord(C->ev_early, ss));
    CU(cudaGetLastError());
  }
  P.ms_part[4] = now_ms() - h6;
  P.ms_host = (float)(now_ms() - h0);
  if (getenv("M2K_DEBUG_TIMING"))
    fprintf(stderr, "[m2k] prepare %.2f ms: order %.2f, parse %.2f, fold %.2f, table %.2f, upload+lists %.2f (%zu files, %zu entries)\n",
            P.ms_host, P.ms_part[0], P.ms_part[1], P.ms_part[2], P.ms_part[3], P.ms_part[4], P.ign_entries.size(), P.folded.size());
  P.ready = true;
  return M2K_OK;
}

extern "C" int m2k_host_prepare_probe(const m2k_tree* tree, const m2k_detect_params* params, uint32_t reps, double* ms_best,
                                      uint32_t* n_rule_files, uint32_t* n_rules) {
  if (!tree || !ms_best) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  // stand-in for the name kernel's list of rule files (not timed): a plain scan of the names
  std::vector<uint32_t> found;
  for (uint32_t e = 0; e < tree->n_entries; e++) {
    uint32_t a = tree->entry_name_off[e], b = tree->entry_name_off[e + 1];
    if (b - a == 10 && !memcmp(tree->names + a, ".m2kignore", 10) && (tree->entry_type[e] & 0xF) != M2K_T_DIR) found.push_back(e);
  }
  IgnorePrep P;
  Staging S;
  S.pinned = false;
  for (int k = 0; k < 6; k++) ms_best[k] = 1e30;
  int rc = M2K_OK;
  std::string err;
  for (uint32_t it = 0; it < (reps ? reps : 1) && rc == M2K_OK; it++) {
    double t0 = now_ms();
    P.reset();
    P.ign_entries = found;
    rc = prepare_host(tree, params, P, S, &err);
    double t1 = now_ms();
    uint32_t n1 = 0, n2 = 0;
    ListBuf b1, b2;
    char** l1 = list_from_arenas(P.arenas, 1, &n1, &b1);
    char** l2 = list_from_arenas(P.arenas, 2, &n2, &b2);
    double t2 = now_ms();
    free(l1); free(l2);
    const double v[6] = {t2 - t0, P.ms_part[0], P.ms_part[1], P.ms_part[2], P.ms_part[3], t2 - t1};
    for (int k = 0; k < 6; k++) ms_best[k] = std::min(ms_best[k], v[k]);
  }
  if (n_rule_files) *n_rule_files = (uint32_t)found.size();
  if (n_rules) *n_rules = (uint32_t)P.folded.size();
  S.release();
  if (rc) return fail(nullptr, rc, "%s", err.c_str());
  return M2K_OK;
}

extern "C" int m2k_prepare(m2k_ctx* ctx, m2k_devtree* dt, const m2k_detect_params* params) {
  if (!ctx || !dt) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  if (!dt->t.scanned) return fail(C, M2K_ERR_ARG, "m2k_prepare before m2k_scan");
  CU(cudaSetDevice(C->device));
  return prepare_ignore(C, dt, params);
}

extern "C" int m2k_finalize(m2k_ctx* ctx, m2k_devtree* dt, const m2k_detect_params* params, m2k_result** out) {
  if (!ctx || !dt || !out) return fail(nullptr, M2K_ERR_ARG, "NULL argument");
  Ctx* C = &ctx->c;
  DevTree* t = &dt->t;
  *out = nullptr;
  if (!t->scanned) return fail(C, M2K_ERR_ARG, "m2k_finalize before m2k_scan");
  CU(cudaSetDevice(C->device));
  cudaStream_t s = C->stream;
  const uint32_t nd = t->n_dirs, nf = t->n_files;
  const uint32_t sh = params ? params->sh_flavor : M2K_SH_BASH;
  // m2k_prepare (under the byte scan) parsed the ignore rules and queued the directory-level phases on
  // the side stream; when it was not called, or was called for the other /bin/sh flavour, that happens now
  if (!dt->prep.ready || dt->prep.sh_flavor != sh) { if (int rc = prepare_ignore(C, dt, params)) return rc; }
  IgnorePrep& P = dt->prep;
  m2k_result* r = dt->pending;
  dt->pending = nullptr;
  struct RG { m2k_result* r; bool keep = false; ~RG() { if (!keep) m2k_result_free(r); } } rg{r};
  r->ms_host_ignore = P.ms_host;
  r->n_services = 0;
  // what the byte scan adds: content bits and main-file ids up the tree, then the remaining arrays
  CU(cudaStreamWaitEvent(s, C->ev_early, 0));
  CU(cudaEventRecord(C->ev[4], s));
  launch_content_up(C, t, s);
  CU(cudaEventRecord(C->ev[5], s));
  // dir_subtree and file_flags are neighbours in the slab and in the host block: one copy
  CU(cudaMemcpyAsync(r->dir_subtree, t->dir_subtree, (size_t)(t->file_flags - (uint8_t*)t->dir_subtree) + nf, cudaMemcpyDeviceToHost, s));
  CU(cudaMemcpyAsync(r->dir_main_file, t->dir_main_sub, 4ull * nd, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamWaitEvent(s, C->ev_feat, 0));  // r->entry_feat
  CU(cudaEventRecord(C->ev[6], s));
  cudaError_t werr = cudaStreamSynchronize(s);
  if (werr != cudaSuccess) return fail(C, M2K_ERR_CUDA, "detection kernels failed: %s", cudaGetErrorString(werr));
  if (t->link.attached) {
    uint32_t aborted = 0;
    CU(cudaMemcpy(&aborted, &t->mailbox->abort, 4, cudaMemcpyDeviceToHost));
    if (aborted) return fail(C, M2K_ERR_CUDA, "multi-GPU pass %u: a rank did not deliver its part within 30 s", t->link.pass);
  }
  if (int rc = collect_scan_timings(C, dt)) return rc;
  float ms_k1 = 0, ms_k2 = 0, ms_d = 0;
  CU(cudaEventElapsedTime(&ms_k1, C->ev_side0, C->ev_side1));  // directory phases on the side stream (under the byte scan)
  CU(cudaEventElapsedTime(&ms_k2, C->ev[4], C->ev[5]));
  CU(cudaEventElapsedTime(&ms_d, C->ev[5], C->ev[6]));
  r->ms_kernels = ms_k1 + ms_k2; r->ms_d2h = ms_d;
  C->phase_ms[4] = ms_k2; C->phase_ms[5] = r->ms_host_ignore; C->phase_ms[7] = ms_d;
  double v0 = now_ms();
  {
    // the directories that became services, in id (= Walk) order: eight states per step, since nearly
    // all of them are something else
    const uint8_t* st = r->dir_state;
    const uint64_t pat = 0x0101010101010101ull * M2K_DIR_SERVICE;
    uint32_t d = 0;
    for (; d + 8 <= nd; d += 8) {
      uint64_t x;
      memcpy(&x, st + d, 8);
      x ^= pat;
      if (((x - 0x0101010101010101ull) & ~x & 0x8080808080808080ull) == 0) continue;  // no byte equals SERVICE
      for (uint32_t k = 0; k < 8; k++) if (st[d + k] == M2K_DIR_SERVICE) r->service_dirs[r->n_services++] = d + k;
    }
    for (; d < nd; d++) if (st[d] == M2K_DIR_SERVICE) r->service_dirs[r->n_services++] = d;
  }
  // (the two ignore lists were filled in by m2k_prepare, under the byte scan)
  C->phase_ms[6] = (float)(now_ms() - v0);
  P.ready = false;
  r->ms_h2d = dt->ms_h2d;
  r->ms_scan_kernel = C->last_scan_ms;
  r->n_kernel_launches = (uint32_t)C->launches;
  rg.keep = true;
  *out = r;
  return M2K_OK;
}

// ---- the whole path: host buffers in, host results out ----------------------------------------
extern "C" int m2k_detect(m2k_ctx* ctx, const m2k_tree* tree, const m2k_detect_params* params, m2k_result** out) {
  if (!ctx || !tree || !out) return fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_detect: NULL argument");
  *out = nullptr;
  const bool content = !(params && params->skip_content);
  if (content && !tree->bytes && tree->content_bytes) return fail(&ctx->c, M2K_ERR_ARG, "m2k_detect needs host bytes (tree->bytes is NULL); use the staged calls for device-filled trees");
  double t0 = now_ms();
  uint64_t l0 = ctx->c.launches;
  m2k_devtree* dt = nullptr;
  int rc = m2k_upload(ctx, tree, 0, content ? tree->n_files : 0, &dt);
  if (rc != M2K_OK) return rc;
  const double t1 = now_ms();
  rc = m2k_scan(ctx, dt, params);
  if (rc == M2K_OK) rc = m2k_finalize(ctx, dt, params, out);
  if (rc != M2K_OK) { m2k_devtree_free(ctx, dt); return rc; }
  const double t2 = now_ms();
  // the result's arrays live in the device tree's pinned buffers: free the device memory now, hand the
  // rest to the result
  cudaStreamSynchronize(ctx->c.stream);
  devtree_release(&dt->t);
  dt->t = DevTree();
  (*out)->owner = dt;
  if (getenv("M2K_DEBUG_TIMING"))
    fprintf(stderr, "[m2k] detect: upload %.2f ms (H2D %.2f), scan + finalize %.2f, release %.2f\n", t1 - t0, dt->ms_h2d, t2 - t1, now_ms() - t2);
  if (rc == M2K_OK) {
    (*out)->ms_total = (float)(now_ms() - t0);
    (*out)->n_kernel_launches = (uint32_t)(ctx->c.launches - l0);
  }
  return rc;
}

// ---- the whole path on several devices, from ONE process ----------------------------------------
// What a single-process host (the Go CLI) calls to use every GPU of the box: one host thread per
// device uploads and scans a shard (ranges of 64 KiB chunks of near-equal cost, large files cut
// between devices), the shards' kernels reduce straight into device 0's arrays over NVLink (peer
// access; m2k_exchange_*), and the calling thread -- rank 0 -- assembles the result.  No collective
// library, no second process.
namespace {
struct Rendezvous {  // all ranks meet `n` at a time; the first error wins
  std::mutex m;
  std::condition_variable cv;
  unsigned n, waiting = 0, generation = 0;
  int rc = M2K_OK;
  std::string err;
  explicit Rendezvous(unsigned n_) : n(n_) {}
  void fail_with(int code, const std::string& msg) { std::lock_guard<std::mutex> lk(m); if (rc == M2K_OK) { rc = code; err = msg; } }
  int meet() {  // returns the common status once everybody has arrived
    std::unique_lock<std::mutex> lk(m);
    const unsigned gen = generation;
    if (++waiting == n) { waiting = 0; generation++; cv.notify_all(); }
    else cv.wait(lk, [&] { return generation != gen; });
    return rc;
  }
};
}  // namespace

extern "C" int m2k_detect_multi(m2k_ctx* const* ctxs, uint32_t n, const m2k_tree* tree, const m2k_detect_params* params, m2k_result** out) {
  if (!ctxs || !n || !ctxs[0] || !tree || !out) return fail(ctxs && n ? &ctxs[0]->c : nullptr, M2K_ERR_ARG, "m2k_detect_multi: NULL argument");
  if (n == 1) return m2k_detect(ctxs[0], tree, params, out);
  Ctx* C = &ctxs[0]->c;
  if (n > kMaxRanks) return fail(C, M2K_ERR_ARG, "m2k_detect_multi: at most %u devices", kMaxRanks);
  for (uint32_t r = 0; r < n; r++) {
    if (!ctxs[r]) return fail(C, M2K_ERR_ARG, "m2k_detect_multi: context %u is NULL", r);
    for (uint32_t q = 0; q < r; q++) if (ctxs[q] == ctxs[r]) return fail(C, M2K_ERR_ARG, "m2k_detect_multi: context %u is given twice (one context per shard)", r);
  }
  *out = nullptr;
  const bool content = !(params && params->skip_content);
  if (!content) return m2k_detect(ctxs[0], tree, params, out);  // names only: nothing to shard
  if (!tree->bytes && tree->content_bytes) return fail(C, M2K_ERR_ARG, "m2k_detect_multi needs host bytes (tree->bytes is NULL)");
  const double t0 = now_ms();
  const uint64_t l0 = C->launches;
  std::vector<uint64_t> cuts(n + 1);
  if (int rc = m2k_plan_shards(tree, n, 64ull << 20, m2k_root_extra_bytes(tree), cuts.data())) return rc;
  std::vector<m2k_devtree*> dts(n, nullptr);
  Rendezvous rv(n);
  auto rank_main = [&](uint32_t r) {
    m2k_ctx* ctx = ctxs[r];
    // phase 1: this rank's shard on its device
    int rc = m2k_upload_chunks(ctx, tree, cuts[r], cuts[r + 1], &dts[r]);
    if (rc != M2K_OK) rv.fail_with(rc, ctx->c.err);
    if (rv.meet() != M2K_OK) return;
    // phase 2: the root opens the exchange, the others attach to its memory
    if (r == 0) {
      m2k_exchange_handle h;
      rc = m2k_exchange_open(ctx, dts[0], n, &h);
      if (rc != M2K_OK) rv.fail_with(rc, ctx->c.err);
    }
    if (rv.meet() != M2K_OK) return;
    if (r != 0) {
      rc = m2k_exchange_attach_local(ctx, dts[r], r, ctxs[0], dts[0]);
      if (rc != M2K_OK) rv.fail_with(rc, ctx->c.err);
    }
    if (rv.meet() != M2K_OK) return;
    // phase 3: scan; attached ranks deliver into the root's arrays and are done
    rc = m2k_scan(ctx, dts[r], params);
    if (rc == M2K_OK && r != 0) rc = m2k_exchange_wait(ctx, dts[r]);
    if (rc != M2K_OK) rv.fail_with(rc, ctx->c.err);
  };
  std::vector<std::thread> threads;
  for (uint32_t r = 1; r < n; r++) threads.emplace_back(rank_main, r);
  rank_main(0);
  int rc = M2K_OK;
  {
    std::lock_guard<std::mutex> lk(rv.m);
    rc = rv.rc;
  }
  m2k_devtree* dt = dts[0];
  if (rc == M2K_OK) rc = m2k_prepare(ctxs[0], dt, params);  // the ignore rules, under the byte scans
  if (rc == M2K_OK) rc = m2k_exchange_collect(ctxs[0], dt);
  for (std::thread& th : threads) th.join();
  {
    std::lock_guard<std::mutex> lk(rv.m);
    if (rc == M2K_OK && rv.rc != M2K_OK) { rc = rv.rc; C->err = rv.err; g_err = rv.err; }
  }
  // a FROM-first file that was cut between devices and has no long line needs its body looked at as a
  // whole (M2K_FILE_UNDECIDED): one pass over that single file on the root
  for (uint32_t r = 1; r < n && rc == M2K_OK; r++) {
    if (cuts[r] == 0 || cuts[r] >= cuts[n] || cuts[r] == cuts[r - 1]) continue;
    uint32_t f = 0, k = 0;
    if (m2k_chunk_locate(tree, cuts[r], &f, &k) != M2K_OK || k == 0) continue;
    uint8_t flags = 0;
    if ((rc = m2k_file_flag(ctxs[0], dt, f, &flags)) != M2K_OK || !(flags & M2K_FILE_UNDECIDED)) continue;
    m2k_devtree* one = nullptr;
    if ((rc = m2k_upload(ctxs[0], tree, f, f + 1, &one)) != M2K_OK) break;
    uint8_t settled = 0;
    rc = m2k_scan(ctxs[0], one, params);
    if (rc == M2K_OK) rc = m2k_file_flag(ctxs[0], one, f, &settled);
    m2k_devtree_free(ctxs[0], one);
    if (rc == M2K_OK) rc = m2k_settle_file(ctxs[0], dt, f, settled & 0xF);
  }
  if (rc == M2K_OK) rc = m2k_finalize(ctxs[0], dt, params, out);
  for (uint32_t r = 1; r < n; r++) if (dts[r]) m2k_devtree_free(ctxs[r], dts[r]);
  if (rc != M2K_OK) { if (dt) m2k_devtree_free(ctxs[0], dt); return rc; }
  cudaSetDevice(C->device);
  cudaStreamSynchronize(C->stream);
  devtree_release(&dt->t);
  dt->t = DevTree();
  (*out)->owner = dt;
  (*out)->ms_total = (float)(now_ms() - t0);
  (*out)->n_kernel_launches = (uint32_t)(C->launches - l0);
  return M2K_OK;
}

// Pack and detect in one call, for a tree that is looked at once: the walk and the numbering as in
// m2k_pack_tree, but the file bytes are never gathered on the host -- a team of threads reads them
// straight into the context's two pinned blocks while the DMA engine drains the other one into the
// device buffer.  *tree_out is the packed tree WITHOUT a bytes blob (names, types, sizes; the
// .m2kignore contents are kept for the ignore rules); release it with m2k_tree_free.
extern "C" int m2k_pack_detect(m2k_ctx* ctx, const char* root_path, const m2k_detect_params* params, m2k_tree** tree_out, m2k_result** out) {
  if (!ctx || !root_path || !tree_out || !out) return fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_pack_detect: NULL argument");
  Ctx* C = &ctx->c;
  *tree_out = nullptr;
  *out = nullptr;
  CU(cudaSetDevice(C->device));
  const bool content = !(params && params->skip_content);
  double t0 = now_ms();
  uint64_t l0 = C->launches;
  PackPlan plan;
  m2k_tree* tree = nullptr;
  std::string err;
  if (int rc = pack_plan(root_path, content ? 1 : 0, false, false, &plan, &tree, &err)) return fail(C, rc, "%s", err.c_str());
  pack_keep_rule_files(&plan, tree);
  m2k_devtree* dt = nullptr;
  int rc = upload_impl(ctx, tree, 0, tree->n_files, 0, M2K_NONE, &dt);
  if (rc != M2K_OK) { tree_release(tree); return rc; }
  auto drop = [&]() { m2k_devtree_free(ctx, dt); tree_release(tree); };
  // stream the bytes: block k is filled by the team (each thread a byte sub-range: files are opened
  // relative to their directory, large files are shared between threads through pread) while block
  // k-1 is on its way to the device
  std::atomic<bool> stale{false};
  {
    if ((rc = ensure_stage(C)) != M2K_OK) { drop(); return rc; }
    const uint64_t total = pack_plan_bytes(&plan);
    cudaStream_t s = C->stream;
    cudaError_t ce = cudaEventRecord(C->ev[0], s);
    int k = 0;
    for (uint64_t lo = 0; lo < total && ce == cudaSuccess; lo += kStageBytes, k ^= 1) {
      const uint64_t len = std::min<uint64_t>(kStageBytes, total - lo);
      ce = cudaEventSynchronize(C->stage_ev[k]);
      if (ce != cudaSuccess) break;
      uint8_t* blk = static_cast<uint8_t*>(C->stage[k]);
      const unsigned nt = std::max(1u, std::min<unsigned>(host_threads(len, 256u << 10), 32));
      run_threads(nt, [&](unsigned tid) {
        uint64_t a = (len * tid / nt) & ~15ull, b = tid + 1 == nt ? len : (len * (tid + 1) / nt) & ~15ull;
        if (b > a && !pack_read_range(&plan, lo + a, lo + b, blk + a)) stale = true;
      });
      ce = cudaMemcpyAsync(dt->t.bytes + lo, blk, len, cudaMemcpyHostToDevice, s);
      if (ce == cudaSuccess) ce = cudaEventRecord(C->stage_ev[k], s);
    }
    if (ce == cudaSuccess) ce = cudaEventRecord(C->ev[1], s);
    if (ce != cudaSuccess) { drop(); return fail(C, M2K_ERR_CUDA, "streaming the tree to the device failed: %s", cudaGetErrorString(ce)); }
  }
  if (stale) {
    // a file shrank between the walk and the read: the layout is stale.  Take the two-step road, whose
    // read pass settles the sizes before anything is uploaded.
    drop();
    m2k_tree* full = nullptr;
    if ((rc = pack_tree_impl(false, root_path, content ? 1 : 0, &full, &err)) != M2K_OK) return fail(C, rc, "%s", err.c_str());
    rc = m2k_detect(ctx, full, params, out);
    if (rc != M2K_OK) { tree_release(full); return rc; }
    *tree_out = full;
    return M2K_OK;
  }
  rc = m2k_scan(ctx, dt, params);
  if (rc == M2K_OK) rc = m2k_finalize(ctx, dt, params, out);
  if (rc != M2K_OK) { drop(); return rc; }
  cudaStreamSynchronize(C->stream);
  float ms_h2d = 0.f;
  if (cudaEventElapsedTime(&ms_h2d, C->ev[0], C->ev[1]) != cudaSuccess) cudaGetLastError();
  devtree_release(&dt->t);
  dt->t = DevTree();
  (*out)->owner = dt;
  (*out)->ms_h2d = ms_h2d;  // here: reading the files and copying them, overlapped
  (*out)->ms_total = (float)(now_ms() - t0);
  (*out)->n_kernel_launches = (uint32_t)(C->launches - l0);
  *tree_out = tree;
  return M2K_OK;
}

// isDockerFile (dockerfile2kube.go:117-144) for ONE file the caller has read: what
// getDockerfileServices does when filepath.Walk hands it a root that is not a directory.  The bytes
// become a one-file tree and take the same kernels as every other file.
extern "C" int m2k_is_dockerfile(m2k_ctx* ctx, const uint8_t* data, uint64_t len, uint32_t* file_flags) {
  if (!ctx || (!data && len) || !file_flags) return fail(ctx ? &ctx->c : nullptr, M2K_ERR_ARG, "m2k_is_dockerfile: NULL argument");
  *file_flags = 0;
  const uint64_t padded = (len + 15) & ~15ull;
  m2k_tree* t = tree_alloc(1, 1, 1, 1, padded + 16, true, false, ".");
  if (!t) return fail(&ctx->c, M2K_ERR_NOMEM, "out of host memory");
  t->dir_parent[0] = M2K_NONE; t->dir_end[0] = 1; t->dir_entry[0] = M2K_NONE;
  t->dir_first_entry[0] = 0; t->dir_first_entry[1] = 1;
  t->entry_dir[0] = 0; t->entry_name_off[0] = 0; t->entry_name_off[1] = 1;
  t->entry_type[0] = M2K_T_REG | (M2K_T_REG << 4); t->entry_child[0] = 0; t->names[0] = 'f';
  t->file_entry[0] = 0; t->file_off[0] = 0; t->file_len[0] = len;
  t->content_bytes = len;
  memset(t->bytes, 0, padded + 16);
  if (len) memcpy(t->bytes, data, len);
  m2k_result* r = nullptr;
  int rc = m2k_detect(ctx, t, nullptr, &r);
  if (rc == M2K_OK) { *file_flags = r->file_flags[0]; m2k_result_free(r); }
  tree_release(t);
  return rc;
}

// ---- host helpers exported for the mirror ----------------------------------------------------
extern "C" int m2k_path_clean(const char* path, char* outbuf, size_t cap) {
  if (!path || !outbuf || !cap) return M2K_ERR_ARG;
  std::string c = path_clean(path);
  if (c.size() + 1 > cap) return M2K_ERR_LIMIT;
  memcpy(outbuf, c.c_str(), c.size() + 1);
  return M2K_OK;
}

extern "C" int m2k_equal_fold(const char* a, size_t na, const char* b, size_t nb) {
  return equal_fold(std::string(a ? a : "", a ? na : 0), std::string(b ? b : "", b ? nb : 0)) ? 1 : 0;
}

extern "C" int m2k_list_entries(const m2k_tree* tree, uint32_t feature_bit, const uint32_t* entry_feat, uint32_t* outv, uint32_t cap) {
  if (!tree || !entry_feat || feature_bit > 31) return M2K_ERR_ARG;
  std::vector<uint32_t> v;
  for (uint32_t e = 0; e < tree->n_entries; e++) if (entry_feat[e] >> feature_bit & 1) v.push_back(e);
  sort_walk_order(tree, &v);
  for (size_t i = 0; outv && i < v.size() && i < cap; i++) outv[i] = v[i];
  return (int)v.size();
}
