/*
 * libm2kdetect -- B200-native source detection for move2kube (C ABI).
 *
 * This is the drop-in boundary for ONE path of the reference: the source-detection pass of
 * `move2kube plan`.  A Go binding (cgo) would declare exactly these
 * entry points; the stub is shown in INTEGRATION.md.  Plain C types only, no callbacks.
 *
 * Reference interfaces replaced (file:line under the reference tree):
 *   Any2KubeTranslator.GetServiceOptions        internal/source/any2kube.go:43-105
 *   Any2KubeTranslator.getIgnorePaths           internal/source/any2kube.go:151-181
 *   common.GetFilesByName / GetFilesByExt       internal/common/utils.go:85-120 / 47-82
 *   common.IsStringPresent (strings.EqualFold)  internal/common/utils.go:308-315
 *   Containerizers.InitContainerizers /
 *     GetContainerizationOptions                internal/containerizer/containerizer.go:56-76
 *   DockerfileContainerizer.GetTargetOptions    internal/containerizer/dockerfilecontainerizer.go:62-83
 *   S2IContainerizer.GetTargetOptions           internal/containerizer/s2icontainerizer.go:63-84
 *   the 18 detect scripts                       internal/assets/{dockerfiles,s2i}/<lang>/m2k*detect.sh
 *   isDockerFile / getDockerfileServices walk   internal/source/dockerfile2kube.go:117-170
 *
 * Error convention: every call returns M2K_OK (0) or a negative m2k_status; the message is
 * kept per context (m2k_last_error) or, when no context exists, per thread
 * (m2k_global_error).  There is no CPU fallback: without a usable CUDA device every compute
 * entry point fails with M2K_ERR_CUDA.
 *
 * Threading: a context is used by one thread at a time (the reference path is
 * single-goroutine).  Calls set the CUDA device themselves.
 */
#ifndef M2KDETECT_H
#define M2KDETECT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define M2K_ABI_VERSION 1
#define M2K_NONE 0xFFFFFFFFu

typedef enum m2k_status {
  M2K_OK = 0,
  M2K_ERR_ARG = -1,         /* bad argument */
  M2K_ERR_CUDA = -2,        /* CUDA runtime error or no device */
  M2K_ERR_IO = -3,          /* root missing / unreadable (GetFilesByName error path) */
  M2K_ERR_NOMEM = -4,
  M2K_ERR_UNSUPPORTED = -5, /* input the packed representation cannot express exactly
                               (e.g. a file name containing a newline) */
  M2K_ERR_LIMIT = -6,       /* a stated size limit was exceeded */
  M2K_ERR_NOTDIR = -7       /* the root exists but Lstat says it is not a directory (a file, or a symlink even
                               to a directory): filepath.Walk visits it alone -- no services (any2kube.go:63-65),
                               and getDockerfileServices asks isDockerFile about the root itself
                               (dockerfile2kube.go:154-167; see m2k_is_dockerfile) */
} m2k_status;

/* ---- entry types: low nibble = lstat type, high nibble = type after following symlinks ---- */
enum { M2K_T_ERR = 0, M2K_T_REG = 1, M2K_T_DIR = 2, M2K_T_LNK = 3, M2K_T_OTHER = 4 };

/* ---- /bin/sh flavour of the python detect scripts ---- */
enum { M2K_SH_BASH = 0, M2K_SH_DASH = 1 };

/* ---- per-entry / per-directory name features (bit numbers) ----
 * F_* : [ -f D/name ]   -- entry exists with that exact basename and is a regular file after
 *                          following symlinks (detect scripts' `[ -f ... ]`)
 * R_* : find -name "*.ext" -- entry of ANY type whose basename ends in ".ext"
 * WAR_OK: first non-dot entry "*.war" of the directory (byte order, LC_ALL=C) exists ([ -e ]) */
enum {
  M2K_F_PIPFILE = 0, M2K_F_BUILD_XML = 1, M2K_F_BUILD_GRADLE = 2, M2K_F_POM_XML = 3,
  M2K_F_PACKAGE_JSON = 4, M2K_F_REQUIREMENTS = 5, M2K_F_SETUP_PY = 6, M2K_F_ENVIRONMENT_YML = 7,
  M2K_F_GEMFILE = 8, M2K_F_GO_MOD = 9,
  M2K_R_GO = 10, M2K_R_PHP = 11, M2K_R_JAVA = 12,
  M2K_N_WAR = 13,        /* entry: non-dot name ending ".war"; directory: WAR_OK */
  M2K_N_IGNOREFILE = 14, /* non-directory named ".m2kignore"          (utils.go:85-120) */
  M2K_N_USER_DF = 15,    /* non-directory named "m2kdfdetect.sh"      (dockerfilecontainerizer.go:50-59) */
  M2K_N_USER_S2I = 16,   /* non-directory named "m2ks2idetect.sh"     (s2icontainerizer.go:51-60) */
  M2K_N_YAML = 17,       /* non-directory whose filepath.Ext is .yml/.yaml (GetFilesByExt callers) */
  M2K_C_MAIN_PY = 18,    /* content: a file whose name ends "py" contains "__main__" */
  M2K_C_DOCKERFILE = 19  /* content: a file isDockerFile() accepts */
};

/* ---- containerizer option bits in dir_opts (= TargetOptions order) ---- */
enum {
  M2K_DF_DJANGO = 0, M2K_DF_GOLANG = 1, M2K_DF_WAR_JBOSS = 2, M2K_DF_WAR_LIBERTY = 3,
  M2K_DF_WAR_TOMCAT = 4, M2K_DF_JAVAANT = 5, M2K_DF_JAVAGRADLE = 6, M2K_DF_JAVAMAVEN = 7,
  M2K_DF_NODEJS = 8, M2K_DF_PHP = 9, M2K_DF_PYTHON = 10, M2K_DF_RUBY = 11,
  M2K_S2I_GOLANG = 12, M2K_S2I_JAVA = 13, M2K_S2I_NODEJS = 14, M2K_S2I_PHP = 15,
  M2K_S2I_PYTHON = 16, M2K_S2I_RUBY = 17,
  M2K_N_OPTS = 18
};
#define M2K_DF_MASK 0x00000FFFu
#define M2K_S2I_MASK 0x0003F000u

/* ---- per-file content flags ---- */
enum {
  M2K_FILE_HAS_MAIN = 1,    /* bytes contain "__main__"              (python detect scripts :21/:22) */
  M2K_FILE_LONG_LINE = 2,   /* some line >= 65536 bytes (bufio.Scanner ErrTooLong) */
  M2K_FILE_FROM_FIRST = 4,  /* first non-ARG instruction is FROM and matches dockerfile2kube.go:131 */
  M2K_FILE_DOCKERFILE = 8,  /* isDockerFile() == true                (dockerfile2kube.go:117-144) */
  M2K_FILE_UNDECIDED = 16   /* only for a file that m2k_upload_chunks spread over several GPUs: HAS_MAIN and
                               LONG_LINE are final, but FROM_FIRST / DOCKERFILE could not be settled from the
                               pieces (the file starts like a Dockerfile and has no long line, so its body
                               needs validating as a whole): scan that one file with m2k_upload(f, f+1) */
};

/* ---- what the walk did with a directory (any2kube.go:58-98) ---- */
enum {
  M2K_DIR_NOT_VISITED = 0,        /* hidden below a pruned directory */
  M2K_DIR_PRECONTAINERIZED = 1,   /* :66-68  SkipDir */
  M2K_DIR_IGNORED_PRUNE = 2,      /* :69-72  SkipDir */
  M2K_DIR_IGNORED_DESCEND = 3,    /* :73     */
  M2K_DIR_NOMATCH_PRUNE = 4,      /* :78-80  SkipDir */
  M2K_DIR_NOMATCH_DESCEND = 5,    /* :81     */
  M2K_DIR_SERVICE = 6,            /* :83-97  services emitted, SkipDir */
  M2K_DIR_WALK_ERROR = 7          /* :59-62  ReadDir failed: warned, not detected */
};

/* ---- the packed tree: struct-of-arrays in (pinned) host memory -------------------------
 * Directories are numbered in DFS pre-order with children in byte order (filepath.Walk's
 * order), so a subtree is the id range [d, dir_end[d]).  Entries (files, directories,
 * symlinks, other) are sorted by (directory id, name).  Files that carry content -- entries
 * whose type after following symlinks is REG -- are numbered in entry order; each file's
 * bytes start 16-byte aligned in `bytes`, zero padded, and at least 16 zero bytes follow
 * the last file.  `bytes` may be NULL when the content lives only on the device
 * (synthetic trees filled by m2k_synth_fill_device). */
typedef struct m2k_tree {
  uint32_t n_dirs, n_entries, n_files, max_depth;
  uint64_t n_bytes;      /* size of the bytes blob including padding */
  uint64_t n_name_bytes; /* size of the names blob */
  uint64_t content_bytes;/* sum of file_len: the algorithmic bytes of one scan */
  const char* root_path; /* as given by the caller (not cleaned), NUL terminated */
  uint32_t* dir_parent;      /* [n_dirs]   parent directory id, root: M2K_NONE */
  uint32_t* dir_end;         /* [n_dirs]   one past the last directory id of the subtree */
  uint32_t* dir_entry;       /* [n_dirs]   this directory's entry index, root: M2K_NONE */
  uint32_t* dir_first_entry; /* [n_dirs+1] entries of d are [dir_first_entry[d], dir_first_entry[d+1]) */
  uint32_t* dir_depth;       /* [n_dirs]   root = 0 */
  uint8_t* dir_flags;        /* [n_dirs]   bit0: ReadDir failed */
  uint32_t* entry_dir;       /* [n_entries] containing directory id (non-decreasing) */
  uint32_t* entry_name_off;  /* [n_entries+1] offsets into names */
  uint8_t* entry_type;       /* [n_entries] lstat type | followed type << 4 */
  uint32_t* entry_child;     /* [n_entries] directory id (DIR), file id (content), else M2K_NONE */
  uint8_t* names;            /* [n_name_bytes] basenames, not NUL terminated */
  uint32_t* file_entry;      /* [n_files] */
  uint64_t* file_off;        /* [n_files] offset into bytes, multiple of 16 */
  uint64_t* file_len;        /* [n_files] */
  uint8_t* bytes;            /* [n_bytes] or NULL */
  void* owner;               /* library bookkeeping */
} m2k_tree;

/* ---- results (host memory, owned by the library, released with m2k_result_free) ----
 * The arrays are pinned buffers the device copies into directly.  A result of m2k_detect stays valid
 * until m2k_result_free.  A result of m2k_finalize borrows the buffers of its device tree: it is valid
 * until the next m2k_scan / m2k_prepare / m2k_finalize on that tree or its m2k_devtree_free, and must
 * be released before or together with the tree.  The two ignore lists live in such buffers too. */
typedef struct m2k_result {
  uint32_t n_dirs, n_files, n_services, n_entries;
  uint32_t* dir_opts;      /* [n_dirs] built-in option bits (M2K_DF_*, M2K_S2I_*) for EVERY directory */
  uint32_t* dir_direct;    /* [n_dirs] features of the directory's own entries */
  uint32_t* dir_subtree;   /* [n_dirs] OR of dir_direct over the subtree (R_* predicates, content bits) */
  uint8_t* dir_state;      /* [n_dirs] M2K_DIR_* */
  uint8_t* dir_ignore;     /* [n_dirs] bit0 in ignoreDirectories, bit1 in ignoreContents, bit2 preContainerized */
  uint32_t* dir_main_file; /* [n_dirs] lowest file id under the directory with M2K_C_MAIN_PY, else M2K_NONE */
  uint8_t* file_flags;     /* [n_files] M2K_FILE_* */
  uint32_t* service_dirs;  /* [n_services] directory ids with state SERVICE, walk order */
  uint32_t* entry_feat;    /* [n_entries] name-feature bits of every entry (M2K_F_*, M2K_R_*, M2K_N_*) */
  uint32_t n_ignore_dirs, n_ignore_contents;
  char** ignore_dirs;      /* cleaned paths, any2kube.go:175-176 */
  char** ignore_contents;  /* any2kube.go:170-173 */
  /* timing of the last run, milliseconds (CUDA events / host clock) */
  float ms_h2d, ms_kernels, ms_scan_kernel, ms_d2h, ms_host_ignore, ms_total;
  uint32_t n_kernel_launches;
  uint32_t _pad;
  void* owner;             /* library bookkeeping */
} m2k_result;

typedef struct m2k_detect_params {
  uint32_t sh_flavor;          /* M2K_SH_BASH (canonical) or M2K_SH_DASH */
  uint32_t n_pre;              /* preContainerizedSourcePaths, any2kube.go:47-54 */
  const char* const* pre_paths;
  uint32_t skip_content;       /* 1: names only (no byte scan): a11/a12 bits stay 0 */
  uint32_t _pad;
} m2k_detect_params;

typedef struct m2k_ctx m2k_ctx;       /* one CUDA device, its stream and scratch */
typedef struct m2k_devtree m2k_devtree; /* a tree (or a shard of its files) resident in HBM */

/* -- lifecycle -- */
int m2k_abi_version(void);
const char* m2k_global_error(void);
int m2k_ctx_create(int device, m2k_ctx** out);
void m2k_ctx_destroy(m2k_ctx* ctx);
const char* m2k_last_error(const m2k_ctx* ctx);
/* a context keeps the two device buffers that grow with a shard (its bytes, the Dockerfile stripe states) and the
 * page-locked host blocks of its results when they are released, and hands them to the next call that fits, so
 * that repeated m2k_detect calls pay neither cudaMalloc / cudaFree of the whole tree nor page-locking each time;
 * m2k_ctx_trim gives the idle ones back to the driver.  A result that is released after its context was
 * destroyed frees what it holds itself (hosts with a garbage collector release in any order); a device tree
 * still needs a live context for m2k_devtree_free. */
int m2k_ctx_trim(m2k_ctx* ctx);
int m2k_device_info(m2k_ctx* ctx, int* sm_count, uint64_t* hbm_bytes, char* name, size_t name_cap);

/* -- host packer: ONE filesystem walk replaces the reference's 4-11 walks
 *    (utils.go:85-120 x3 per Any2Kube call, any2kube.go:58, dockerfile2kube.go:154) -- */
int m2k_pack_tree(m2k_ctx* ctx, const char* root_path, int read_content, m2k_tree** out);
void m2k_tree_free(m2k_ctx* ctx, m2k_tree* tree);
/* pinned host memory for callers that stage file bytes themselves (NULL on failure, see m2k_last_error) */
void* m2k_pinned_alloc(m2k_ctx* ctx, size_t bytes);
void m2k_pinned_free(m2k_ctx* ctx, void* p);

/* -- synthetic trees of BASELINE.md section 4 (configs B, C, D), deterministic in `seed` -- */
typedef struct m2k_synth_params {
  uint64_t seed;
  uint32_t n_files, n_dirs, max_depth, n_ignore_files;
  uint64_t total_bytes;   /* target content size */
  uint32_t n_huge;        /* config D: number of multi-GiB files */
  uint32_t _pad;
  uint64_t huge_bytes;    /* size of each */
  uint32_t tiny_min, tiny_max; /* config D: tiny file size range (0,0 = log-normal sizes) */
  uint32_t n_large;       /* config D: a second group of big files */
  uint32_t _pad2;
  uint64_t large_bytes;
} m2k_synth_params;
int m2k_synth_tree(m2k_ctx* ctx, const m2k_synth_params* p, int host_bytes, m2k_tree** out);
/* bytes of file ids [file_begin, file_end) generated on the host into dst (file_off relative to
 * file_off[file_begin]); the SAME generator the device fill uses */
int m2k_synth_fill_host(const m2k_tree* tree, uint64_t seed, uint32_t file_begin, uint32_t file_end,
                        uint8_t* dst, int n_threads);

/* -- whole path, host buffers in, host results out (H2D + kernels + D2H): the call a plugin makes -- */
int m2k_detect(m2k_ctx* ctx, const m2k_tree* tree, const m2k_detect_params* params, m2k_result** out);
void m2k_result_free(m2k_result* r);
/* -- the same on SEVERAL devices from one process (the reference's caller is one process: any2kube.go:43-46):
 * ctxs[0..n) are contexts on distinct devices (or, for testing, on the same one); one host thread per context
 * uploads and scans a shard (ranges of 64 KiB chunks of near-equal cost, files of >= 64 MiB may be cut between
 * devices), the shards' kernels reduce straight into ctxs[0]'s arrays over NVLink (m2k_exchange_* with peer access;
 * no collective library), and the calling thread assembles the result on ctxs[0].  Same result as m2k_detect. -- */
int m2k_detect_multi(m2k_ctx* const* ctxs, uint32_t n, const m2k_tree* tree, const m2k_detect_params* params, m2k_result** out);
/* -- the same for a tree that is looked at once, straight from the file system: the walk and numbering
 * of m2k_pack_tree, but the file bytes are never gathered on the host -- host threads read them into
 * the context's two pinned blocks while the DMA engine drains the other one.  *tree_out is the packed
 * tree without a bytes blob (names, types, sizes, and the contents of the .m2kignore files); release
 * it with m2k_tree_free(ctx, tree) after m2k_result_free. -- */
int m2k_pack_detect(m2k_ctx* ctx, const char* root_path, const m2k_detect_params* params, m2k_tree** tree_out, m2k_result** out);

/* -- isDockerFile (dockerfile2kube.go:117-144) for one file the caller has read, through the same kernels:
 * what getDockerfileServices needs when its root is not a directory (M2K_ERR_NOTDIR).  *file_flags gets the
 * M2K_FILE_* bits; M2K_FILE_DOCKERFILE is isDockerFile's answer. -- */
int m2k_is_dockerfile(m2k_ctx* ctx, const uint8_t* data, uint64_t len, uint32_t* file_flags);

/* -- staged form (bench, multi-GPU): a shard = file ids [file_begin, file_end) -- */
int m2k_upload(m2k_ctx* ctx, const m2k_tree* tree, uint32_t file_begin, uint32_t file_end, m2k_devtree** out);
/* the same shard again, host to device, into the buffers m2k_upload / m2k_upload_chunks allocated (no allocation:
 * for a caller that keeps its device buffers between calls); `tree` must have the shape of the original */
int m2k_reupload(m2k_ctx* ctx, m2k_devtree* dt, const m2k_tree* tree);
/* -- a shard as a range of 64 KiB chunks, so that one multi-GB file can be spread over several GPUs
 * (SURVEY.md section 8e).  Chunks are numbered over the whole tree: files in id order, each
 * ceil(file_len / 65536) chunks (m2k_chunk_count = the total).  A shard may begin and end inside a
 * file; it then also holds the 16 bytes on the far side of the cut.  m2k_plan_shards picks
 * n_ranks + 1 cut points of near-equal cost (bytes + a fixed part per chunk) that fall inside a file only when the file has at least
 * `split_min_bytes` bytes (smaller files go whole to one side).  The parts of a cut file are
 * summarised per shard (m2k_pieces: 64 bytes, device to device), gathered in rank order and stitched
 * by m2k_merge_pieces on the rank that finalises; see M2K_FILE_UNDECIDED for the one case left open. */
uint64_t m2k_chunk_count(const m2k_tree* tree);
int m2k_plan_shards(const m2k_tree* tree, uint32_t n_ranks, uint64_t split_min_bytes, uint64_t root_extra_bytes,
                    uint64_t* chunk_cuts /* [n_ranks+1] */);
/* root_extra_bytes: shard 0 is made lighter by this much, because its rank also runs the
 * directory-level phases and assembles the result (0 = equal shards); m2k_root_extra_bytes is the
 * library's own estimate for a tree, from timings on B200 */
uint64_t m2k_root_extra_bytes(const m2k_tree* tree);
int m2k_upload_chunks(m2k_ctx* ctx, const m2k_tree* tree, uint64_t chunk_begin, uint64_t chunk_end, m2k_devtree** out);
uint64_t m2k_devtree_content_bytes(const m2k_devtree* dt); /* content bytes resident in this shard (the algorithmic bytes of its scan) */
int m2k_pieces(m2k_ctx* ctx, m2k_devtree* dt, void* d_dst /* 64 bytes */);
int m2k_merge_pieces(m2k_ctx* ctx, m2k_devtree* dt, const void* d_all_pieces /* n_ranks x 64 bytes, rank order */, uint32_t n_ranks);
/* the same exchange without any host synchronisation: this rank's contribution copied into the
 * caller's device buffers (n_dirs uint32, n_dirs uint32, 64 bytes, n_files bytes), and the gathered
 * data (rank-major arrays, the summed flags) merged from the caller's device buffers, all on the
 * library's stream (m2k_stream).  The caller orders its collectives' stream against that stream with
 * events and leaves the buffers alone until m2k_finalize has returned. */
int m2k_export_async(m2k_ctx* ctx, m2k_devtree* dt, uint32_t* d_content, uint32_t* d_main, void* d_pieces, uint8_t* d_flags);
int m2k_import_async(m2k_ctx* ctx, m2k_devtree* dt, const uint32_t* d_all_content, const uint32_t* d_all_main,
                     const void* d_all_pieces, const uint8_t* d_flags, uint32_t n_ranks);
/* -- the exchange of a sharded pass over PEER MEMORY (NVLink / NVSwitch), no collective library in the data
 * path ("fused peer atomicOr into rank 0's buffer").  One rank -- the root, rank 0, the one
 * that calls m2k_prepare / m2k_finalize -- opens the exchange on its device tree; every other rank attaches its own
 * device tree (same packed tree, another shard) to the root's memory: through the 192-byte handle when it lives
 * in another process (CUDA IPC; ship the bytes any way you like, e.g. one broadcast at set-up), or directly when
 * it lives in the same process on another (or the same) device.  From then on m2k_scan on an attached rank skips
 * the name kernel and, once the root has cleared its arrays for the pass, its per-file kernels OR / MIN their
 * per-directory results straight into the ROOT's arrays; its own range of the per-file flags and its two piece
 * summaries follow, then a flag in the root's mailbox.  The root calls m2k_exchange_collect between its m2k_scan
 * and m2k_finalize: it waits (on the device) for every rank's flag and stitches the files that were cut between
 * ranks.  All ranks must call m2k_scan the same number of times.  A rank that never arrives makes the pass fail
 * after 30 s (m2k_finalize returns M2K_ERR_CUDA) instead of hanging the GPU. -- */
typedef struct m2k_exchange_handle { uint8_t opaque[192]; } m2k_exchange_handle;
int m2k_exchange_open(m2k_ctx* ctx, m2k_devtree* root_dt, uint32_t n_ranks, m2k_exchange_handle* out);
int m2k_exchange_attach(m2k_ctx* ctx, m2k_devtree* dt, uint32_t rank, const m2k_exchange_handle* handle);
int m2k_exchange_attach_local(m2k_ctx* ctx, m2k_devtree* dt, uint32_t rank, m2k_ctx* root_ctx, m2k_devtree* root_dt);
int m2k_exchange_collect(m2k_ctx* ctx, m2k_devtree* root_dt);
/* an attached rank: block (without spinning) until its part of the current pass has been delivered */
int m2k_exchange_wait(m2k_ctx* ctx, m2k_devtree* dt);
/* ... or only until the pass BEFORE the last m2k_scan has been delivered: a rank that serves pass after pass may
 * queue the next one at once (its writes into the root's memory wait on the device for the root's flag), so its
 * next byte scan starts the moment its GPU is free */
int m2k_exchange_wait_previous(m2k_ctx* ctx, m2k_devtree* dt);
/* which file holds global chunk `chunk`, and which of its chunks that is (a cut at chunk_in_file != 0 splits the file) */
int m2k_chunk_locate(const m2k_tree* tree, uint64_t chunk, uint32_t* file, uint32_t* chunk_in_file);
/* one file's flags on the device, read back (synchronises); and, before m2k_finalize, the flags a
 * separate pass over that one file found for a file left M2K_FILE_UNDECIDED (also records the
 * Dockerfile bit of its directory) */
int m2k_file_flag(m2k_ctx* ctx, m2k_devtree* dt, uint32_t file, uint8_t* out);
int m2k_settle_file(m2k_ctx* ctx, m2k_devtree* dt, uint32_t file, uint32_t flags);
int m2k_synth_fill_device(m2k_ctx* ctx, m2k_devtree* dt, uint64_t seed);
/* phase A: name features + byte scan + Dockerfile check over the resident shard */
int m2k_scan(m2k_ctx* ctx, m2k_devtree* dt, const m2k_detect_params* params);
/* per-directory partial arrays a multi-GPU run gathers, copied device-to-device into the caller's
 * buffers (n_dirs uint32 each): content mask (OR-reducible) and lowest main-file id (MIN-reducible) */
int m2k_partials(m2k_ctx* ctx, m2k_devtree* dt, uint32_t* d_dst_content, uint32_t* d_dst_main);
/* OR / MIN `n_ranks` gathered partial arrays (device memory, rank-major) into this shard's own */
int m2k_merge_partials(m2k_ctx* ctx, m2k_devtree* dt, const uint32_t* d_all_content, const uint32_t* d_all_main, uint32_t n_ranks);
/* per-file flags of a sharded run, copied device-to-device out of / into the device tree ([n_files] bytes):
 * each rank writes only its own file range, so a byte-wise sum across ranks is the union */
int m2k_file_flags_get(m2k_ctx* ctx, m2k_devtree* dt, uint8_t* d_dst);
int m2k_file_flags_set(m2k_ctx* ctx, m2k_devtree* dt, const uint8_t* d_src);
/* optional, between m2k_scan and m2k_finalize: the host half of getIgnorePaths (any2kube.go:151-181) --
 * parse the .m2kignore files the name kernel found and upload the lookup table -- while the byte scan
 * is still running on the GPU.  m2k_finalize does it itself when this was not called. */
int m2k_prepare(m2k_ctx* ctx, m2k_devtree* dt, const m2k_detect_params* params);
/* diagnostic, no counterpart in the reference and no device needed: runs only the HOST half of
 * m2k_prepare (order the rule files like filepath.Walk, parse, Clean/Join, fold, build the table, the
 * two lists) `reps` times on a tree whose rule files are readable on the host, and reports the best
 * time of each part in ms_best[6] = {all, order, parse, fold, table, lists}.  It produces no
 * detection result: it exists so that this serial stretch of a multi-GPU pass can be timed alone. */
int m2k_host_prepare_probe(const m2k_tree* tree, const m2k_detect_params* params, uint32_t reps, double* ms_best,
                           uint32_t* n_rule_files, uint32_t* n_rules);
/* phase C: subtree OR, predicate combine, ignore join, prune; copies results to the host */
int m2k_finalize(m2k_ctx* ctx, m2k_devtree* dt, const m2k_detect_params* params, m2k_result** out);
void m2k_devtree_free(m2k_ctx* ctx, m2k_devtree* dt);
/* device time of the byte-scan kernel alone in the last m2k_scan, and launches issued so far */
int m2k_scan_kernel_ms(m2k_ctx* ctx, float* ms);
/* milliseconds of the last m2k_scan + m2k_finalize by phase: name features, byte scan, per-file
 * combine, isDockerFile, finalize kernels (device); ignore parsing, hit verification (host); D2H */
int m2k_phase_ms(m2k_ctx* ctx, float* out, int cap);
uint64_t m2k_launch_count(const m2k_ctx* ctx);
void* m2k_stream(m2k_ctx* ctx); /* cudaStream_t the library launches on */

/* diagnostic, no counterpart in the reference: a library built with -DM2K_CHECKED=1 counts, instead of making it,
 * every kernel access that falls outside the buffer it belongs to (scan loads outside the shard, record / list /
 * stripe indices past their tables).  out[0..8) receives the counters and clears them; returns 1 for a checked
 * build, 0 for the product build (whose counters stay zero), or a negative m2k_status. */
int m2k_debug_violations(m2k_ctx* ctx, uint64_t* out);

/* -- the byte scan without any GPU allocation (no counterpart in the reference) --
 * For hosts that cannot call cudaMalloc where they call this library: inside work they are
 * capturing or replaying, under a device-memory arena that is handed out once and then locked, or
 * under a fixed memory quota.  The caller owns every buffer; the scan's only scratch, the per-chunk
 * record table, is a fixed 2 MiB array in the library image (held by every GPU that loads the
 * library until exit), which limits one call to 262144 chunks of 64 KiB (16 GiB of content): more
 * is refused with M2K_ERR_LIMIT.  All pointers are DEVICE pointers on `device`:
 *   d_bytes        file contents; each file starts at a multiple of 16, is zero padded to 16 bytes,
 *                  and at least 16 zero bytes follow the last file
 *   d_file_off/len [n_files] offset into d_bytes / length in bytes
 *   d_chunk_first  [n_files+1] exclusive prefix sum of ceil(len / 65536); n_chunks MUST equal
 *                  d_chunk_first[n_files] (the host cannot read it; a prefix that runs past n_chunks is cut off there)
 *   d_file_flags   [n_files] out: M2K_FILE_HAS_MAIN | M2K_FILE_LONG_LINE
 * Work is queued on `stream` (a cudaStream_t, NULL = default stream) and not waited for; the caller's
 * current device is the same after the call as before it.  The first
 * call on a device looks up the table's device address: make it before any stream capture.  Calls
 * share the table; run them one after the other (overlapping calls race on it).  Nothing in the
 * table needs clearing between calls: each call writes every record it later reads; the one word that
 * must start at zero (the counter the scan draws its chunks from, next to the table) is cleared on
 * `stream` by every call.
 * Error text: m2k_global_error(). */
int m2k_scan_flags_noalloc(int device, void* stream, const uint8_t* d_bytes, const uint64_t* d_file_off,
                           const uint64_t* d_file_len, const uint32_t* d_chunk_first, uint32_t n_files,
                           uint32_t n_chunks, uint8_t* d_file_flags);

/* -- translate-time re-detection without re-running anything: the bytes the accepting
 * detect script of built-in option `option` (M2K_DF_* / M2K_S2I_*) prints on stdout for directory `dir` -- the JSON
 * DockerfileContainerizer.GetContainer / S2IContainerizer.GetContainer re-run the script for
 * (dockerfilecontainerizer.go:105-129, s2icontainerizer.go:97-115): constants, the directory's first *.war entry, or
 * the main script the byte scan found (`main_script_rel_path` / `app_file`: the lowest file id under the directory;
 * the reference's answer is the first path `grep -R` prints, which depends on readdir order).  Returns the number
 * of bytes written (without the terminating NUL; buf == NULL only asks for the size), M2K_ERR_ARG when the option
 * does not accept the directory (its script prints nothing and exits 1), M2K_ERR_LIMIT when cap is too small.
 * Host code, no device needed; `res` must come from a pass with content (skip_content = 0) for the python options. -- */
int m2k_option_json(const m2k_tree* tree, const m2k_result* res, uint32_t dir, uint32_t option, char* buf, size_t cap);

/* -- host-side path helpers the mirror needs (Go stdlib semantics) -- */
int m2k_path_clean(const char* path, char* out, size_t cap);              /* filepath.Clean */
int m2k_equal_fold(const char* a, size_t na, const char* b, size_t nb);   /* strings.EqualFold */
/* filepath.Walk-order listing of non-directory entries selected by feature bit (GetFilesByName /
 * GetFilesByExt replacement): writes entry indices, returns count */
int m2k_list_entries(const m2k_tree* tree, uint32_t feature_bit, const uint32_t* entry_feat, uint32_t* out, uint32_t cap);

#ifdef __cplusplus
}
#endif
#endif /* M2KDETECT_H */
