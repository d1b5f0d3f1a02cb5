/* TEST INFRASTRUCTURE (oracle/): CPU restatement of the reference's source-detection pass.
 * See m2k_oracle.c.  Loaded through ctypes by tests/, __graft_entry__.smoke() and bench.py's
 * CPU-baseline legs only. */
#ifndef M2K_ORACLE_H
#define M2K_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#include "../include/m2kdetect.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct m2ko_result {
  uint32_t n_dirs, n_files, n_services, _pad;
  uint32_t* dir_opts;      /* [n_dirs] built-in options of EVERY directory */
  uint8_t* dir_state;      /* [n_dirs] M2K_DIR_* */
  uint8_t* dir_ignore;     /* [n_dirs] (visited directories only) bit0 ignoreDirectories, bit1 ignoreContents, bit2 pre */
  uint32_t* dir_main_file; /* [n_dirs] */
  uint32_t* service_dirs;  /* [n_services] */
  uint8_t* file_flags;     /* [n_files] M2K_FILE_* */
  uint32_t n_ignore_dirs, n_ignore_contents;
  char** ignore_dirs;
  char** ignore_contents;
} m2ko_result;

int m2ko_pack_fs(const char* root, int read_content, m2k_tree** out);
void m2ko_tree_free(m2k_tree* t);
int m2ko_detect(const m2k_tree* t, int sh_flavor, uint32_t n_pre, const char* const* pre_paths, int scan_content,
                int n_threads, m2ko_result** out);
void m2ko_result_free(m2ko_result* r);
uint32_t m2ko_dir_options(const m2k_tree* t, uint32_t d, int sh_flavor);
uint32_t m2ko_files_by_name(const m2k_tree* t, const char* name, uint32_t* out, uint32_t cap);
char* m2ko_dir_path(const m2k_tree* t, uint32_t d);
uint32_t m2ko_dockerfile_flags(const uint8_t* data, uint64_t len);
uint32_t m2ko_file_flags(const uint8_t* data, uint64_t len);
int m2ko_scan_files(const m2k_tree* t, uint32_t f0, uint32_t f1, uint8_t* out, int n_threads);
int m2ko_equal_fold(const char* a, size_t na, const char* b, size_t nb);
char* m2ko_clean(const char* p, size_t n);
char* m2ko_join(const char* a, const char* b);
void m2ko_free(void* p);
#ifdef __cplusplus
}
#endif
#endif
