/* The source-detection pass from plain C, through the C ABI alone (include/m2kdetect.h): what a cgo / JNI / ctypes
 * binding does, without any of them.  One line per service directory and accepted option, in Walk order --
 *
 *     <relative dir>\t<option number>\t<what the option's detect script prints on stdout>
 *
 * -- i.e. Any2KubeTranslator.GetServiceOptions (internal/source/any2kube.go:43-103) followed by the re-detection of
 * GetContainer (internal/containerizer/dockerfilecontainerizer.go:105-129, s2icontainerizer.go:97-115), then one line
 * "dockerfile\t<relative path>" per file isDockerFile accepts (internal/source/dockerfile2kube.go:154-167).
 *
 *     gcc -std=c11 -Iinclude examples/detect.c -Lmove2kube_b200 -lm2kdetect -Wl,-rpath,$PWD/move2kube_b200 -o detect
 *     ./detect /path/to/source [device]
 *
 * tests/test_host.py compiles and links it (no GPU needed for that); tests/test_gpu_mirror.py runs it on a B200 and
 * compares its lines with the Python mirror's.  There is no CPU path: without a device m2k_ctx_create fails and says why. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "m2kdetect.h"

/* path of directory d relative to the root ("." for the root itself) */
static void dir_rel_path(const m2k_tree* t, uint32_t d, char* out, size_t cap) {
  uint32_t chain[4096];
  size_t n = 0, pos = 0;
  for (uint32_t x = d; x != M2K_NONE && t->dir_parent[x] != M2K_NONE && n < 4096; x = t->dir_parent[x]) chain[n++] = x;
  if (n == 0) { snprintf(out, cap, "."); return; }
  while (n-- > 0) {
    const uint32_t e = t->dir_entry[chain[n]];
    const uint32_t a = t->entry_name_off[e], b = t->entry_name_off[e + 1];
    if (pos + (b - a) + 2 > cap) break;
    if (pos) out[pos++] = '/';
    memcpy(out + pos, t->names + a, b - a);
    pos += b - a;
  }
  out[pos] = 0;
}

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s SOURCE_DIR [device]\n", argv[0]); return 2; }
  m2k_ctx* ctx = NULL;
  int rc = m2k_ctx_create(argc > 2 ? atoi(argv[2]) : 0, &ctx);
  if (rc != M2K_OK) { fprintf(stderr, "m2k_ctx_create: %s\n", m2k_global_error()); return 1; }
  m2k_detect_params params;
  memset(&params, 0, sizeof params);
  params.sh_flavor = M2K_SH_BASH;
  m2k_tree* tree = NULL;
  m2k_result* res = NULL;
  rc = m2k_pack_detect(ctx, argv[1], &params, &tree, &res); /* walk + read + copy + kernels + results, one call */
  if (rc == M2K_ERR_NOTDIR) { /* filepath.Walk on a root that is not a directory: no services (any2kube.go:58) */
    m2k_ctx_destroy(ctx);
    return 0;
  }
  if (rc != M2K_OK) { fprintf(stderr, "m2k_pack_detect: %s\n", m2k_last_error(ctx)); m2k_ctx_destroy(ctx); return 1; }
  char path[8192], json[8192];
  for (uint32_t k = 0; k < res->n_services; k++) {
    const uint32_t d = res->service_dirs[k];
    dir_rel_path(tree, d, path, sizeof path);
    for (uint32_t opt = 0; opt < M2K_N_OPTS; opt++) {
      if (!((res->dir_opts[d] >> opt) & 1u)) continue;
      const int n = m2k_option_json(tree, res, d, opt, json, sizeof json);
      if (n < 0) { fprintf(stderr, "m2k_option_json: %s\n", m2k_global_error()); continue; }
      while (n > 0 && json[strlen(json) - 1] == '\n') json[strlen(json) - 1] = 0; /* `echo` ends its line, `printf` does not */
      printf("%s\t%u\t%s\n", path, opt, json);
    }
  }
  for (uint32_t f = 0; f < res->n_files; f++) {
    if (!(res->file_flags[f] & M2K_FILE_DOCKERFILE)) continue;
    const uint32_t e = tree->file_entry[f];
    dir_rel_path(tree, tree->entry_dir[e], path, sizeof path);
    printf("dockerfile\t%s/%.*s\n", path, (int)(tree->entry_name_off[e + 1] - tree->entry_name_off[e]), (const char*)tree->names + tree->entry_name_off[e]);
  }
  m2k_result_free(res);
  m2k_tree_free(ctx, tree);
  m2k_ctx_destroy(ctx);
  return 0;
}
