/*
 * TEST INFRASTRUCTURE (oracle/).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may build, load or call this file.  The product
 * (move2kube_b200/) never links it and has no CPU path of its own.
 *
 * Plain-C restatement of the reference's source-detection pass, written to be obviously
 * faithful rather than fast: every predicate re-walks the subtree like the reference's
 * `find`, ignore look-ups are linear EqualFold scans like common.IsStringPresent.
 * Pinning: tests/test_oracle.py checks it against the JSON fixtures in tests/golden/, which were produced
 * by running the reference's own detect scripts (oracle/script_oracle.py) on the reference's
 * fixtures.  The Dockerfile half (m2ko_is_dockerfile) restates moby/buildkit v0.7.2
 * frontend/dockerfile/parser (go.mod:22), whose source is NOT under /root/reference and which
 * no reference test exercises: that half is "parity unpinned".  oracle/dockerfile_re.py derives the
 * same function a third time by another method (a regex engine, literal string building); the CPU
 * tests hold the three against each other.
 *
 * Each function cites the reference lines it follows (paths under /root/reference).
 */
#define _GNU_SOURCE
#include <ctype.h>
#include <dirent.h>
#include <errno.h>
#include <fcntl.h>
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../include/m2kdetect.h" /* data layout (m2k_tree) and bit numbers only */
#include "casefold_table.h"
#include "m2k_oracle.h"

/* ------------------------------------------------------------------------------------------
 * Go stdlib restatements [external: Go 1.15 stdlib, not under /root/reference]
 * ------------------------------------------------------------------------------------------ */

/* utf8.DecodeRune: invalid or truncated -> (U+FFFD, 1) */
static uint32_t decode_rune(const uint8_t* s, size_t n, size_t* w) {
  uint8_t c = s[0];
  *w = 1;
  if (c < 0x80) return c;
  int need;
  uint8_t lo = 0x80, hi = 0xBF;
  uint32_t r;
  if (c >= 0xC2 && c <= 0xDF) { need = 1; r = c & 0x1F; }
  else if (c >= 0xE0 && c <= 0xEF) { need = 2; r = c & 0x0F; if (c == 0xE0) lo = 0xA0; if (c == 0xED) hi = 0x9F; }
  else if (c >= 0xF0 && c <= 0xF4) { need = 3; r = c & 0x07; if (c == 0xF0) lo = 0x90; if (c == 0xF4) hi = 0x8F; }
  else return 0xFFFD;
  if ((size_t)need >= n) return 0xFFFD;
  if (s[1] < lo || s[1] > hi) return 0xFFFD;
  r = (r << 6) | (s[1] & 0x3F);
  for (int k = 2; k <= need; k++) {
    if (s[k] < 0x80 || s[k] > 0xBF) return 0xFFFD;
    r = (r << 6) | (s[k] & 0x3F);
  }
  *w = (size_t)need + 1;
  return r;
}

/* unicode.IsSpace */
static int is_space_rune(uint32_t r) {
  switch (r) {
    case 0x09: case 0x0A: case 0x0B: case 0x0C: case 0x0D: case 0x20: case 0x85: case 0xA0:
    case 0x1680: case 0x2028: case 0x2029: case 0x202F: case 0x205F: case 0x3000:
      return 1;
  }
  return r >= 0x2000 && r <= 0x200A;
}

/* representative of the rune's simple-folding class; EqualFold(a,b) <=> equal key sequences */
static uint32_t fold_key(uint32_t r) {
  if (r < 0x80) return (r >= 'A' && r <= 'Z') ? r + 32 : r;
  int lo = 0, hi = M2K_CASEFOLD_N - 1;
  while (lo <= hi) {
    int mid = (lo + hi) / 2;
    uint32_t v = m2k_casefold_from[mid];
    if (v == r) return m2k_casefold_to[mid];
    if (v < r) lo = mid + 1; else hi = mid - 1;
  }
  return r;
}

/* strings.EqualFold -- internal/common/utils.go:310 */
int m2ko_equal_fold(const char* a, size_t na, const char* b, size_t nb) {
  size_t i = 0, j = 0;
  while (i < na && j < nb) {
    size_t wa, wb;
    uint32_t ra = decode_rune((const uint8_t*)a + i, na - i, &wa);
    uint32_t rb = decode_rune((const uint8_t*)b + j, nb - j, &wb);
    if (fold_key(ra) != fold_key(rb)) return 0;
    i += wa; j += wb;
  }
  return i == na && j == nb;
}

/* filepath.Clean (Unix).  Returns malloc'd string. */
char* m2ko_clean(const char* p, size_t n) {
  if (n == 0) return strdup(".");
  int rooted = p[0] == '/';
  char* out = (char*)malloc(n + 2);
  size_t w = 0, r = 0, dotdot = 0;
  if (rooted) { out[w++] = '/'; r = 1; dotdot = 1; }
  while (r < n) {
    if (p[r] == '/') { r++; }
    else if (p[r] == '.' && (r + 1 == n || p[r + 1] == '/')) { r++; }
    else if (p[r] == '.' && p[r + 1] == '.' && (r + 2 == n || p[r + 2] == '/')) {
      r += 2;
      if (w > dotdot) { w--; while (w > dotdot && out[w] != '/') w--; }
      else if (!rooted) { if (w > 0) out[w++] = '/'; out[w++] = '.'; out[w++] = '.'; dotdot = w; }
    } else {
      if ((rooted && w != 1) || (!rooted && w != 0)) out[w++] = '/';
      for (; r < n && p[r] != '/'; r++) out[w++] = p[r];
    }
  }
  if (w == 0) { out[w++] = '.'; }
  out[w] = 0;
  return out;
}

/* filepath.Join(a, b): empty elements dropped, then Clean */
char* m2ko_join(const char* a, const char* b) {
  size_t na = strlen(a), nb = strlen(b);
  if (na == 0 && nb == 0) return strdup("");
  char* t = (char*)malloc(na + nb + 2);
  size_t w = 0;
  if (na) { memcpy(t, a, na); w = na; }
  if (na && nb) t[w++] = '/';
  if (nb) { memcpy(t + w, b, nb); w += nb; }
  char* c = m2ko_clean(t, w);
  free(t);
  return c;
}

/* filepath.Dir */
static char* go_dir(const char* p) {
  const char* s = strrchr(p, '/');
  size_t n = s ? (size_t)(s - p) + 1 : 0;
  return m2ko_clean(p, n);
}

/* strings.TrimSpace over [*b, *e) */
static void trim_space(const uint8_t** b, const uint8_t** e) {
  const uint8_t* s = *b; const uint8_t* t = *e;
  while (s < t) { size_t w; uint32_t r = decode_rune(s, (size_t)(t - s), &w); if (!is_space_rune(r)) break; s += w; }
  while (t > s) {
    /* utf8.DecodeLastRune */
    const uint8_t* st = t - 1;
    uint32_t r; size_t w = 1;
    if (*st < 0x80) r = *st;
    else {
      const uint8_t* lim = (t - s >= 4) ? t - 4 : s;
      const uint8_t* q = st;
      while (q > lim && (*q & 0xC0) == 0x80) q--;
      size_t ww; r = decode_rune(q, (size_t)(t - q), &ww);
      if (q + ww != t) { r = 0xFFFD; w = 1; } else w = ww;
    }
    if (!is_space_rune(r)) break;
    t -= w;
  }
  *b = s; *e = t;
}

/* ------------------------------------------------------------------------------------------
 * own filesystem walk -> m2k_tree (independent of the product's packer)
 * filepath.Walk semantics: Lstat, children sorted by byte order, symlinks never followed for
 * descent.  [external: path/filepath, os]
 * ------------------------------------------------------------------------------------------ */
typedef struct { void* p; size_t n, cap, esz; } vec;
static void* vpush(vec* v) {
  if (v->n == v->cap) { v->cap = v->cap ? v->cap * 2 : 64; v->p = realloc(v->p, v->cap * v->esz); }
  return (char*)v->p + (v->n++) * v->esz;
}
static int cmp_names(const void* a, const void* b) { return strcmp(*(char* const*)a, *(char* const*)b); }

typedef struct {
  vec dir_parent, dir_end, dir_entry, dir_first, dir_depth, dir_flags;
  vec entry_dir, entry_name_off, entry_type, entry_child, names;
  vec file_entry, file_off, file_len, bytes;
  int read_content; int newline_name; uint32_t max_depth;
} builder;

static uint8_t mode_type(mode_t m) { return S_ISREG(m) ? M2K_T_REG : S_ISDIR(m) ? M2K_T_DIR : S_ISLNK(m) ? M2K_T_LNK : M2K_T_OTHER; }

typedef struct { uint32_t dir; size_t nn; char** names; uint8_t* types; uint32_t* child; } dirrec;

/* Builds directories in DFS pre-order.  Entries must end up sorted by (dir id, name): since a
 * directory's entries are appended when the directory is numbered (pre-order) that holds
 * by construction if we emit all entries of a directory before descending. */
static void build_dir(builder* B, const char* path, uint32_t parent, uint32_t my_entry, uint32_t depth, vec* todo_first) {
  (void)todo_first;
  uint32_t id = (uint32_t)B->dir_parent.n;
  *(uint32_t*)vpush(&B->dir_parent) = parent;
  *(uint32_t*)vpush(&B->dir_end) = 0;
  *(uint32_t*)vpush(&B->dir_entry) = my_entry;
  *(uint32_t*)vpush(&B->dir_depth) = depth;
  *(uint8_t*)vpush(&B->dir_flags) = 0;
  if (depth > B->max_depth) B->max_depth = depth;
  /* entries of this dir are appended later, in a second pass (see m2ko_pack_fs) */
  DIR* dp = opendir(path);
  char** names = NULL; size_t nn = 0, cap = 0;
  if (!dp) { ((uint8_t*)B->dir_flags.p)[id] = 1; }
  else {
    struct dirent* de;
    while ((de = readdir(dp)) != NULL) {
      if (!strcmp(de->d_name, ".") || !strcmp(de->d_name, "..")) continue;
      if (nn == cap) { cap = cap ? cap * 2 : 16; names = (char**)realloc(names, cap * sizeof(char*)); }
      names[nn++] = strdup(de->d_name);
    }
    closedir(dp);
    qsort(names, nn, sizeof(char*), cmp_names);
  }
  /* stash the sorted names on the builder: we number child directories now (pre-order) and
   * record (dir id -> names, types) for the entry pass */
  size_t plen = strlen(path);
  dirrec* rec = (dirrec*)vpush(todo_first);
  rec->dir = id; rec->nn = nn; rec->names = names;
  rec->types = (uint8_t*)malloc(nn ? nn : 1); rec->child = (uint32_t*)malloc((nn ? nn : 1) * sizeof(uint32_t));
  size_t rec_index = todo_first->n - 1;
  for (size_t i = 0; i < nn; i++) {
    size_t cl = plen + 1 + strlen(names[i]) + 1;
    char* child = (char*)malloc(cl);
    snprintf(child, cl, "%s/%s", path, names[i]);
    struct stat st, st2;
    uint8_t lt = M2K_T_ERR, ft = M2K_T_ERR;
    if (lstat(child, &st) == 0) {
      lt = mode_type(st.st_mode);
      if (lt == M2K_T_LNK) { if (stat(child, &st2) == 0) ft = mode_type(st2.st_mode); }
      else ft = lt;
    }
    if (strchr(names[i], '\n')) B->newline_name = 1;
    uint32_t ch = M2K_NONE;
    if (lt == M2K_T_DIR) {
      ch = (uint32_t)B->dir_parent.n;
      build_dir(B, child, id, 0 /* fixed up in entry pass */, depth + 1, todo_first);
      rec = (dirrec*)todo_first->p + rec_index; /* vector may have moved */
    }
    rec->types[i] = (uint8_t)(lt | (ft << 4));
    rec->child[i] = ch;
    free(child);
  }
  ((uint32_t*)B->dir_end.p)[id] = (uint32_t)B->dir_parent.n;
}

static m2k_tree* tree_from_builder(builder* B, const char* root);

int m2ko_pack_fs(const char* root, int read_content, m2k_tree** out) {
  struct stat st;
  *out = NULL;
  if (stat(root, &st) != 0) return errno == ENOENT ? M2K_ERR_IO : M2K_ERR_IO;
  builder B; memset(&B, 0, sizeof B);
  B.dir_parent.esz = B.dir_end.esz = B.dir_entry.esz = B.dir_first.esz = B.dir_depth.esz = 4; B.dir_flags.esz = 1;
  B.entry_dir.esz = B.entry_name_off.esz = B.entry_child.esz = 4; B.entry_type.esz = 1; B.names.esz = 1;
  B.file_entry.esz = 4; B.file_off.esz = B.file_len.esz = 8; B.bytes.esz = 1;
  B.read_content = read_content;
  vec recs; memset(&recs, 0, sizeof recs); recs.esz = sizeof(dirrec);
  build_dir(&B, root, M2K_NONE, M2K_NONE, 0, &recs);
  /* entry pass: recs are in pre-order = dir id order */
  uint32_t n_dirs = (uint32_t)B.dir_parent.n;
  char** dirpath = (char**)calloc(n_dirs, sizeof(char*));
  dirpath[0] = strdup(root);
  for (size_t k = 0; k < recs.n; k++) {
    dirrec* r = (dirrec*)recs.p + k;
    *(uint32_t*)vpush(&B.dir_first) = (uint32_t)B.entry_dir.n;
    for (size_t i = 0; i < r->nn; i++) {
      uint32_t e = (uint32_t)B.entry_dir.n;
      *(uint32_t*)vpush(&B.entry_dir) = r->dir;
      *(uint32_t*)vpush(&B.entry_name_off) = (uint32_t)B.names.n;
      size_t nl = strlen(r->names[i]);
      for (size_t c = 0; c < nl; c++) *(uint8_t*)vpush(&B.names) = (uint8_t)r->names[i][c];
      *(uint8_t*)vpush(&B.entry_type) = r->types[i];
      uint32_t ch = r->child[i];
      size_t cl = strlen(dirpath[r->dir]) + 1 + nl + 1;
      char* full = (char*)malloc(cl);
      snprintf(full, cl, "%s/%s", dirpath[r->dir], r->names[i]);
      if ((r->types[i] & 0xF) == M2K_T_DIR) {
        ((uint32_t*)B.dir_entry.p)[ch] = e;
        dirpath[ch] = full; full = NULL;
      } else if ((r->types[i] >> 4) == M2K_T_REG) {
        ch = (uint32_t)B.file_entry.n;
        *(uint32_t*)vpush(&B.file_entry) = e;
        uint64_t off = (uint64_t)B.bytes.n, len = 0;
        if (read_content) {
          int fd = open(full, O_RDONLY);
          if (fd >= 0) {
            uint8_t buf[65536]; ssize_t g;
            while ((g = read(fd, buf, sizeof buf)) > 0) {
              for (ssize_t c = 0; c < g; c++) *(uint8_t*)vpush(&B.bytes) = buf[c];
              len += (uint64_t)g;
            }
            close(fd);
          }
          while (B.bytes.n % 16) *(uint8_t*)vpush(&B.bytes) = 0;
        }
        *(uint64_t*)vpush(&B.file_off) = off;
        *(uint64_t*)vpush(&B.file_len) = len;
      }
      *(uint32_t*)vpush(&B.entry_child) = ch;
      free(full);
      free(r->names[i]);
    }
    free(r->names); free(r->types); free(r->child);
  }
  *(uint32_t*)vpush(&B.dir_first) = (uint32_t)B.entry_dir.n;
  *(uint32_t*)vpush(&B.entry_name_off) = (uint32_t)B.names.n;
  for (int k = 0; k < 16; k++) *(uint8_t*)vpush(&B.bytes) = 0;
  for (uint32_t d = 0; d < n_dirs; d++) free(dirpath[d]);
  free(dirpath); free(recs.p);
  m2k_tree* t = tree_from_builder(&B, root);
  *out = t;
  return B.newline_name ? M2K_ERR_UNSUPPORTED : M2K_OK;
}

static m2k_tree* tree_from_builder(builder* B, const char* root) {
  m2k_tree* t = (m2k_tree*)calloc(1, sizeof *t);
  t->n_dirs = (uint32_t)B->dir_parent.n; t->n_entries = (uint32_t)B->entry_dir.n; t->n_files = (uint32_t)B->file_entry.n;
  t->max_depth = B->max_depth;
  t->n_bytes = B->bytes.n; t->n_name_bytes = B->names.n;
  t->root_path = strdup(root);
  t->dir_parent = (uint32_t*)B->dir_parent.p; t->dir_end = (uint32_t*)B->dir_end.p; t->dir_entry = (uint32_t*)B->dir_entry.p;
  t->dir_first_entry = (uint32_t*)B->dir_first.p; t->dir_depth = (uint32_t*)B->dir_depth.p; t->dir_flags = (uint8_t*)B->dir_flags.p;
  t->entry_dir = (uint32_t*)B->entry_dir.p; t->entry_name_off = (uint32_t*)B->entry_name_off.p; t->entry_type = (uint8_t*)B->entry_type.p;
  t->entry_child = (uint32_t*)B->entry_child.p; t->names = (uint8_t*)B->names.p;
  t->file_entry = (uint32_t*)B->file_entry.p; t->file_off = (uint64_t*)B->file_off.p; t->file_len = (uint64_t*)B->file_len.p;
  t->bytes = (uint8_t*)B->bytes.p;
  uint64_t cb = 0;
  for (uint32_t f = 0; f < t->n_files; f++) cb += t->file_len[f];
  t->content_bytes = cb;
  return t;
}

void m2ko_tree_free(m2k_tree* t) {
  if (!t) return;
  free((void*)t->root_path);
  free(t->dir_parent); free(t->dir_end); free(t->dir_entry); free(t->dir_first_entry); free(t->dir_depth); free(t->dir_flags);
  free(t->entry_dir); free(t->entry_name_off); free(t->entry_type); free(t->entry_child); free(t->names);
  free(t->file_entry); free(t->file_off); free(t->file_len); free(t->bytes);
  free(t);
}

/* ------------------------------------------------------------------------------------------
 * tree helpers
 * ------------------------------------------------------------------------------------------ */
static const uint8_t* ename(const m2k_tree* t, uint32_t e, size_t* n) {
  *n = t->entry_name_off[e + 1] - t->entry_name_off[e];
  return t->names + t->entry_name_off[e];
}

/* Walk path of directory d: root as given; children = filepath.Join(parent, name) */
char* m2ko_dir_path(const m2k_tree* t, uint32_t d) {
  if (d == 0) return strdup(t->root_path);
  char* pp = m2ko_dir_path(t, t->dir_parent[d]);
  size_t n; const uint8_t* nm = ename(t, t->dir_entry[d], &n);
  char* name = (char*)malloc(n + 1); memcpy(name, nm, n); name[n] = 0;
  char* j = m2ko_join(pp, name);
  free(pp); free(name);
  return j;
}

static int name_is(const m2k_tree* t, uint32_t e, const char* s) {
  size_t n; const uint8_t* nm = ename(t, e, &n);
  return n == strlen(s) && memcmp(nm, s, n) == 0;
}
static int name_ends(const m2k_tree* t, uint32_t e, const char* suf) {
  size_t n; const uint8_t* nm = ename(t, e, &n);
  size_t k = strlen(suf);
  return n >= k && memcmp(nm + n - k, suf, k) == 0;
}

/* F(name): `[ -f "D/name" ]` -- exists and is a regular file after following symlinks.
 * e.g. internal/assets/dockerfiles/nodejs/m2kdfdetect.sh:16 */
static int pred_F(const m2k_tree* t, uint32_t d, const char* name) {
  for (uint32_t e = t->dir_first_entry[d]; e < t->dir_first_entry[d + 1]; e++)
    if (name_is(t, e, name)) return (t->entry_type[e] >> 4) == M2K_T_REG;
  return 0;
}

/* R(ext): `find "D"/. -name "*.ext" -print -quit | wc -l` == 1 -- an entry of any type strictly
 * below D whose basename ends in .ext; find does not follow symlinks.
 * internal/assets/dockerfiles/golang/m2kdfdetect.sh:16 (names with newlines are refused by the packer) */
static int pred_R(const m2k_tree* t, uint32_t d, const char* ext) {
  for (uint32_t e = t->dir_first_entry[d]; e < t->dir_first_entry[d + 1]; e++) {
    if (name_ends(t, e, ext)) return 1;
    if ((t->entry_type[e] & 0xF) == M2K_T_DIR && pred_R(t, t->entry_child[e], ext)) return 1;
  }
  return 0;
}

/* W: `main "D/"*.war` with `[ ! -e "$1" ] && exit 1` -- the FIRST glob match (no dotfiles, byte
 * collation under LC_ALL=C) must exist after following symlinks.
 * internal/assets/dockerfiles/java-war-tomcat/m2kdfdetect.sh:20-26 */
static int pred_W(const m2k_tree* t, uint32_t d) {
  for (uint32_t e = t->dir_first_entry[d]; e < t->dir_first_entry[d + 1]; e++) {
    size_t n; const uint8_t* nm = ename(t, e, &n);
    if (n >= 4 && nm[0] != '.' && memcmp(nm + n - 4, ".war", 4) == 0) return (t->entry_type[e] >> 4) != M2K_T_ERR;
  }
  return 0;
}

/* the 18 built-in predicates; script lines cited per bit */
uint32_t m2ko_dir_options(const m2k_tree* t, uint32_t d, int sh_flavor) {
  uint32_t m = 0;
  int py = sh_flavor == M2K_SH_BASH &&
           (pred_F(t, d, "requirements.txt") || pred_F(t, d, "setup.py") || pred_F(t, d, "environment.yml") || pred_F(t, d, "Pipfile"));
  if (pred_F(t, d, "Pipfile")) m |= 1u << M2K_DF_DJANGO;          /* dockerfiles/django/m2kdfdetect.sh:18-20 */
  if (pred_R(t, d, ".go")) m |= 1u << M2K_DF_GOLANG;              /* dockerfiles/golang/m2kdfdetect.sh:16-20 */
  if (pred_W(t, d)) m |= (1u << M2K_DF_WAR_JBOSS) | (1u << M2K_DF_WAR_LIBERTY) | (1u << M2K_DF_WAR_TOMCAT); /* java-war-*:20-26 */
  if (pred_F(t, d, "build.xml")) m |= 1u << M2K_DF_JAVAANT;       /* javaant:16-18 */
  if (pred_F(t, d, "build.gradle")) m |= 1u << M2K_DF_JAVAGRADLE; /* javagradle:16-18 */
  if (pred_F(t, d, "pom.xml")) m |= 1u << M2K_DF_JAVAMAVEN;       /* javamaven:16-18 */
  if (pred_F(t, d, "package.json")) m |= 1u << M2K_DF_NODEJS;     /* nodejs:16-18 */
  if (pred_R(t, d, ".php")) m |= 1u << M2K_DF_PHP;                /* php:18-22 */
  if (py) m |= 1u << M2K_DF_PYTHON;                               /* python:17-28 (array syntax: bash only) */
  if (pred_F(t, d, "Gemfile")) m |= 1u << M2K_DF_RUBY;            /* ruby:15-17 */
  if (pred_F(t, d, "go.mod") || pred_R(t, d, ".go")) m |= 1u << M2K_S2I_GOLANG; /* s2i/golang:19-29 */
  if (!pred_F(t, d, "build.gradle") && !pred_F(t, d, "build.xml") && (pred_F(t, d, "pom.xml") || pred_R(t, d, ".java")))
    m |= 1u << M2K_S2I_JAVA;                                      /* s2i/java:21-41 */
  if (pred_F(t, d, "package.json")) m |= 1u << M2K_S2I_NODEJS;    /* s2i/nodejs:16-18 */
  if (pred_R(t, d, ".php")) m |= 1u << M2K_S2I_PHP;               /* s2i/php:19-23 */
  if (py) m |= 1u << M2K_S2I_PYTHON;                              /* s2i/python:17-29 */
  if (pred_F(t, d, "Gemfile")) m |= 1u << M2K_S2I_RUBY;           /* s2i/ruby:19-21 */
  return m;
}

/* ------------------------------------------------------------------------------------------
 * GetFilesByName in Walk order -- internal/common/utils.go:85-120
 * ------------------------------------------------------------------------------------------ */
static void files_by_name_rec(const m2k_tree* t, uint32_t d, const char* name, vec* out) {
  if (t->dir_flags[d] & 1) return;
  for (uint32_t e = t->dir_first_entry[d]; e < t->dir_first_entry[d + 1]; e++) {
    uint8_t lt = t->entry_type[e] & 0xF;
    if (lt == M2K_T_ERR) continue; /* lstat failed: callback gets err, skipped (:98-101) */
    if (lt == M2K_T_DIR) files_by_name_rec(t, t->entry_child[e], name, out);
    else if (name_is(t, e, name)) *(uint32_t*)vpush(out) = e;
  }
}
uint32_t m2ko_files_by_name(const m2k_tree* t, const char* name, uint32_t* out, uint32_t cap) {
  vec v; memset(&v, 0, sizeof v); v.esz = 4;
  files_by_name_rec(t, 0, name, &v);
  uint32_t n = (uint32_t)v.n;
  for (uint32_t i = 0; i < n && i < cap; i++) out[i] = ((uint32_t*)v.p)[i];
  free(v.p);
  return n;
}

/* ------------------------------------------------------------------------------------------
 * getIgnorePaths -- internal/source/any2kube.go:151-181
 * ------------------------------------------------------------------------------------------ */
typedef struct { char** s; size_t n, cap; } strlist;
static void sl_push(strlist* l, char* s) {
  if (l->n == l->cap) { l->cap = l->cap ? l->cap * 2 : 16; l->s = (char**)realloc(l->s, l->cap * sizeof(char*)); }
  l->s[l->n++] = s;
}
static void sl_free(strlist* l) { for (size_t i = 0; i < l->n; i++) free(l->s[i]); free(l->s); }

/* IsStringPresent -- internal/common/utils.go:308-315 (linear EqualFold scan) */
static int is_string_present(const strlist* l, const char* v) {
  size_t nv = strlen(v);
  for (size_t i = 0; i < l->n; i++)
    if (m2ko_equal_fold(l->s[i], strlen(l->s[i]), v, nv)) return 1;
  return 0;
}

static void get_ignore_paths(const m2k_tree* t, strlist* ign_dirs, strlist* ign_contents) {
  uint32_t n = m2ko_files_by_name(t, ".m2kignore", NULL, 0);
  if (!n) return;
  uint32_t* es = (uint32_t*)malloc(n * sizeof(uint32_t));
  m2ko_files_by_name(t, ".m2kignore", es, n);
  for (uint32_t k = 0; k < n; k++) {
    uint32_t e = es[k];
    if ((t->entry_type[e] >> 4) != M2K_T_REG || !t->bytes) continue; /* open/read fails: no lines (:158-162) */
    uint32_t f = t->entry_child[e];
    const uint8_t* data = t->bytes + t->file_off[f];
    uint64_t len = t->file_len[f];
    char* dpath = m2ko_dir_path(t, t->entry_dir[e]);
    char* fpath = m2ko_join(dpath, ".m2kignore");
    char* dir = go_dir(fpath);
    uint64_t pos = 0;
    while (pos < len) { /* bufio.Scanner / ScanLines, 64 KiB token cap [external] */
      const uint8_t* nl = (const uint8_t*)memchr(data + pos, '\n', len - pos);
      uint64_t end = nl ? (uint64_t)(nl - data) : len;
      if (end - pos >= 65536) break; /* ErrTooLong: scanning stops silently */
      const uint8_t* b = data + pos; const uint8_t* en = data + end;
      if (en > b && en[-1] == '\r') en--;
      trim_space(&b, &en);
      size_t ll = (size_t)(en - b);
      int star = ll > 0 && b[ll - 1] == '*';
      if (star) ll--;
      char* line = (char*)malloc(ll + 1); memcpy(line, b, ll); line[ll] = 0;
      /* a NUL inside the line truncates the C string; Go keeps it -- such paths cannot name a directory anyway */
      char* p = m2ko_join(dir, line);
      sl_push(star ? ign_contents : ign_dirs, p);
      free(line);
      pos = end + 1;
    }
    free(dpath); free(fpath); free(dir);
  }
  free(es);
}

/* ------------------------------------------------------------------------------------------
 * content signature A: `grep -lRe "__main__"`  (dockerfiles/python/m2kdfdetect.sh:21)
 * ------------------------------------------------------------------------------------------ */
static int has_main(const uint8_t* p, uint64_t n) { return n >= 8 && memmem(p, n, "__main__", 8) != NULL; }

/* ------------------------------------------------------------------------------------------
 * content signature B: isDockerFile -- internal/source/dockerfile2kube.go:117-144 over
 * moby/buildkit v0.7.2 frontend/dockerfile/parser.Parse [external, restated from the published
 * algorithm; parity unpinned]
 * ------------------------------------------------------------------------------------------ */
typedef struct { char* p; size_t n, cap; } sbuf;
static void sb_add(sbuf* s, const char* d, size_t n) {
  if (s->n + n + 1 > s->cap) { s->cap = (s->n + n + 1) * 2; s->p = (char*)realloc(s->p, s->cap); }
  memcpy(s->p + s->n, d, n); s->n += n; s->p[s->n] = 0;
}

static int is_ascii_ws(uint8_t c) { return c == '\t' || c == '\v' || c == '\f' || c == '\r' || c == ' '; } /* tokenWhitespace */
static int is_re_s(uint8_t c) { return c == '\t' || c == '\n' || c == '\f' || c == '\r' || c == ' '; }     /* RE2 \s */

typedef struct { uint32_t escape; int processing_complete, escape_seen; } directive;

/* possibleParserDirective: `^#[ \t]*escape[ \t]*=[ \t]*(?P<escapechar>.).*$` on the lower-cased line.
 * Returns 0 ok, -1 error. */
static int possible_directive(directive* d, const char* line, size_t n) {
  if (d->processing_complete) return 0;
  size_t i = 0;
  int m = 0;
  if (n > 0 && line[0] == '#') {
    i = 1;
    while (i < n && (line[i] == ' ' || line[i] == '\t')) i++;
    if (i + 6 <= n && strncasecmp(line + i, "escape", 6) == 0) {
      i += 6;
      while (i < n && (line[i] == ' ' || line[i] == '\t')) i++;
      if (i < n && line[i] == '=') {
        i++;
        size_t k = i;
        while (i < n && (line[i] == ' ' || line[i] == '\t')) i++;
        /* `.` needs one character.  The greedy [ \t]* hands a blank back when nothing else follows,
         * so "escape= " still matches, with a blank as the (invalid) escape character. */
        if (i < n) m = 1;
        else if (i > k) { m = 1; i = i - 1; }
      }
    }
  }
  if (m) {
    if (d->escape_seen) return -1; /* only one escape parser directive can be used */
    d->escape_seen = 1;
    /* strings.ToLower runs first: '`' and '\\' are unchanged by it */
    uint8_t c = (uint8_t)line[i];
    if (c != '`' && c != '\\') return -1; /* invalid ESCAPE */
    d->escape = c;
    return 0;
  }
  d->processing_complete = 1;
  return 0;
}

/* trimContinuationCharacter: regexp `\<esc>[ \t]*$` */
static int trim_continuation(const directive* d, sbuf* line) {
  size_t n = line->n;
  while (n > 0 && (line->p[n - 1] == ' ' || line->p[n - 1] == '\t')) n--;
  if (n > 0 && (uint8_t)line->p[n - 1] == d->escape) { line->n = n - 1; line->p[line->n] = 0; return 0; /* not end of line */ }
  return 1;
}

/* parseWords: returns number of words; *first_has_eq, *all_have_eq */
static size_t parse_words(const char* rest, size_t n, uint32_t esc, int* first_has_eq, int* all_have_eq) {
  enum { inSpaces, inWord, inQuote } phase = inSpaces;
  size_t words = 0, wlen = 0; int weq = 0, blankOK = 0; uint32_t quote = 0, ch = 0; size_t w = 0;
  *first_has_eq = 0; *all_have_eq = 1;
#define END_WORD() do { if (blankOK || wlen > 0) { if (words == 0) *first_has_eq = weq; if (!weq) *all_have_eq = 0; words++; } } while (0)
  for (size_t pos = 0; pos <= n; pos += w) {
    if (pos != n) ch = decode_rune((const uint8_t*)rest + pos, n - pos, &w);
    if (phase == inSpaces) {
      if (pos == n) break;
      if (is_space_rune(ch)) continue;
      phase = inWord;
    }
    if ((phase == inWord || phase == inQuote) && pos == n) { END_WORD(); break; }
    if (phase == inWord) {
      if (is_space_rune(ch)) { phase = inSpaces; END_WORD(); wlen = 0; weq = 0; blankOK = 0; continue; }
      if (ch == '\'' || ch == '"') { quote = ch; blankOK = 1; phase = inQuote; }
      if (ch == esc) {
        if (pos + w == n) continue; /* escape token at end of line: skipped */
        wlen += w; pos += w;
        ch = decode_rune((const uint8_t*)rest + pos, n - pos, &w);
      }
      wlen += w; if (ch == '=') weq = 1;
      continue;
    }
    if (phase == inQuote) {
      if (ch == quote) phase = inWord;
      if (ch == esc && quote != '\'') {
        if (pos + w == n) { phase = inWord; continue; }
        pos += w; wlen += w;
        ch = decode_rune((const uint8_t*)rest + pos, n - pos, &w);
      }
      wlen += w; if (ch == '=') weq = 1;
    }
  }
#undef END_WORD
  return words;
}

/* extractBuilderFlags: returns offset of the remaining part (n if none).  Byte-wise, with
 * unicode.IsSpace applied to the BYTE value (so 0x85 and 0xA0 count as space). */
static size_t skip_builder_flags(const char* line, size_t n) {
  enum { inSpaces, inWord, inQuote } phase = inSpaces;
  size_t wlen = 0; int is_dashdash = 0; uint8_t quote = 0, ch = 0;
  /* track whether word == "--": word bytes are appended; compare incrementally */
  char w2[3] = {0, 0, 0};
  for (size_t pos = 0; pos <= n; pos++) {
    if (pos != n) ch = (uint8_t)line[pos];
    if (phase == inSpaces) {
      if (pos == n) break;
      if (is_space_rune(ch)) continue;
      if (ch != '-' || pos + 1 == n || line[pos + 1] != '-') return pos;
      phase = inWord; wlen = 0;
    }
    if ((phase == inWord || phase == inQuote) && pos == n) break;
    is_dashdash = (wlen == 2 && w2[0] == '-' && w2[1] == '-');
    if (phase == inWord) {
      if (is_space_rune(ch)) {
        phase = inSpaces;
        if (is_dashdash) return pos;
        wlen = 0;
        continue;
      }
      if (ch == '\'' || ch == '"') { quote = ch; phase = inQuote; continue; }
      if (ch == '\\') {
        if (pos + 1 == n) continue;
        pos++; ch = (uint8_t)line[pos];
      }
      if (wlen < 2) w2[wlen] = (char)ch;
      wlen++;
      continue;
    }
    if (phase == inQuote) {
      if (ch == quote) { phase = inWord; continue; }
      if (ch == '\\') {
        if (pos + 1 == n) { phase = inWord; continue; }
        pos++; ch = (uint8_t)line[pos];
      }
      if (wlen < 2) w2[wlen] = (char)ch;
      wlen++;
    }
  }
  return n;
}

/* strings.ToLower on the command word, compared with an ASCII keyword: the only non-ASCII runes
 * that lower-case into ASCII are U+0130 -> 'i' and U+212A -> 'k'. */
static int cmd_is(const char* w, size_t n, const char* kw) {
  size_t i = 0, k = 0, kn = strlen(kw);
  while (i < n) {
    size_t cw; uint32_t r = decode_rune((const uint8_t*)w + i, n - i, &cw);
    if (r >= 'A' && r <= 'Z') r += 32; else if (r == 0x130) r = 'i'; else if (r == 0x212A) r = 'k';
    if (k >= kn || r != (uint8_t)kw[k]) return 0;
    i += cw; k++;
  }
  return k == kn;
}

/* parseJSON's json.NewDecoder(...).Decode(&[]interface{}) on s[0..n), s[0] == '[': the FIRST JSON value of the
 * stream by encoding/json's scanner (what follows the closing bracket is not looked at).  Returns 1 when the array is
 * well formed, every number in it (at any depth) is in float64 range (d.convertNumber -> strconv.ParseFloat fails on
 * overflow only) and some top-level element is not a string: errDockerfileNotStringArray, the one JSON outcome that
 * makes parseMaybeJSON / parseMaybeJSONToList -- and with them Parse -- fail.  Anything else (syntax error, unclosed
 * array, nesting beyond Go's 10000, number out of range, all strings) returns 0: the argument is then taken as a plain
 * string.  [external] Go 1.15 encoding/json scanner.go, decode.go; buildkit v0.7.2 line_parsers.go. */
static int json_array_not_all_strings(const uint8_t* s, size_t n) {
  enum { OBJ_KEY, OBJ_VALUE, ARR_VALUE };
  enum { BEGIN_VALUE, BEGIN_VALUE_OR_EMPTY, BEGIN_STRING_OR_EMPTY, BEGIN_STRING, IN_STRING, IN_ESC, NEG, N1, N0, DOT, DOT0, E, ESIGN, E0, END_VALUE };
  size_t cap = 64, depth = 0, i = 0, num_start = 0;
  uint8_t* st = (uint8_t*)malloc(cap);
  int step = BEGIN_VALUE, non_string = 0, out_of_range = 0, result = 0;
#define JFAIL() do { free(st); return 0; } while (0)
#define JPUSH(kind) do { if (depth == cap) { cap *= 2; st = (uint8_t*)realloc(st, cap); } st[depth++] = (uint8_t)(kind); if (depth > 10000) JFAIL(); } while (0)
#define JNUM_END() do { char tmp[64]; char* lit = (i - num_start < sizeof tmp) ? tmp : (char*)malloc(i - num_start + 1); \
    memcpy(lit, s + num_start, i - num_start); lit[i - num_start] = 0; errno = 0; double v = strtod(lit, NULL); \
    if (v == HUGE_VAL || v == -HUGE_VAL) out_of_range = 1; \
    if (lit != tmp) free(lit); } while (0)
  for (;;) {
    if (i >= n) JFAIL(); /* io.ErrUnexpectedEOF */
    const uint8_t c = s[i];
    const int sp = c == ' ' || c == '\t' || c == '\r' || c == '\n';
    int redo = 0;
    switch (step) {
      case BEGIN_VALUE: case BEGIN_VALUE_OR_EMPTY: case BEGIN_STRING_OR_EMPTY: case BEGIN_STRING:
        if (sp) break;
        if (step == BEGIN_VALUE_OR_EMPTY && c == ']') { step = END_VALUE; redo = 1; break; }
        if (step == BEGIN_STRING_OR_EMPTY && c == '}') { st[depth - 1] = OBJ_VALUE; step = END_VALUE; redo = 1; break; }
        if (step == BEGIN_STRING_OR_EMPTY || step == BEGIN_STRING) { if (c != '"') JFAIL(); step = IN_STRING; break; }
        if (depth == 1 && c != '"') non_string = 1;
        if (c == '{') { JPUSH(OBJ_KEY); step = BEGIN_STRING_OR_EMPTY; }
        else if (c == '[') { JPUSH(ARR_VALUE); step = BEGIN_VALUE_OR_EMPTY; }
        else if (c == '"') step = IN_STRING;
        else if (c == '-') { step = NEG; num_start = i; }
        else if (c == '0') { step = N0; num_start = i; }
        else if (c >= '1' && c <= '9') { step = N1; num_start = i; }
        else if (c == 't' || c == 'f' || c == 'n') {
          const char* w = c == 't' ? "true" : c == 'f' ? "false" : "null";
          const size_t wn = strlen(w);
          if (n - i < wn || memcmp(s + i, w, wn) != 0) JFAIL();
          i += wn - 1; step = END_VALUE;
        } else JFAIL();
        break;
      case IN_STRING:
        if (c == '"') step = END_VALUE; else if (c == '\\') step = IN_ESC; else if (c < 0x20) JFAIL();
        break;
      case IN_ESC:
        if (c == 'b' || c == 'f' || c == 'n' || c == 'r' || c == 't' || c == '\\' || c == '/' || c == '"') step = IN_STRING;
        else if (c == 'u') {
          if (n - i < 5) JFAIL();
          for (int k = 1; k <= 4; k++) if (!isxdigit(s[i + k])) JFAIL();
          i += 4; step = IN_STRING;
        } else JFAIL();
        break;
      case NEG: if (c == '0') step = N0; else if (c >= '1' && c <= '9') step = N1; else JFAIL(); break;
      case N1: if (!(c >= '0' && c <= '9')) { step = N0; redo = 1; } break;
      case N0: if (c == '.') step = DOT; else if (c == 'e' || c == 'E') step = E; else { JNUM_END(); step = END_VALUE; redo = 1; } break;
      case DOT: if (!(c >= '0' && c <= '9')) JFAIL(); step = DOT0; break;
      case DOT0: if (c >= '0' && c <= '9') break; if (c == 'e' || c == 'E') step = E; else { JNUM_END(); step = END_VALUE; redo = 1; } break;
      case E: if (c == '+' || c == '-') step = ESIGN; else { step = ESIGN; redo = 1; } break;
      case ESIGN: if (!(c >= '0' && c <= '9')) JFAIL(); step = E0; break;
      case E0: if (!(c >= '0' && c <= '9')) { JNUM_END(); step = END_VALUE; redo = 1; } break;
      case END_VALUE:
        if (sp) break;
        if (st[depth - 1] == OBJ_KEY) { if (c != ':') JFAIL(); st[depth - 1] = OBJ_VALUE; step = BEGIN_VALUE; }
        else if (st[depth - 1] == OBJ_VALUE) {
          if (c == ',') { st[depth - 1] = OBJ_KEY; step = BEGIN_STRING; } else if (c == '}') depth--; else JFAIL();
        } else {
          if (c == ',') step = BEGIN_VALUE;
          else if (c == ']') { if (--depth == 0) { result = non_string && !out_of_range; free(st); return result; } }
          else JFAIL();
        }
        break;
    }
    if (!redo) i++;
  }
#undef JFAIL
#undef JPUSH
#undef JNUM_END
}

/* parseMaybeJSON / parseMaybeJSONToList on [b, e): -1 when they fail */
static int maybe_json(const uint8_t* b, const uint8_t* e) {
  while (b < e) { size_t w; uint32_t r = decode_rune(b, (size_t)(e - b), &w); if (!is_space_rune(r)) break; b += w; } /* parseJSON: TrimLeftFunc */
  if (b == e || *b != '[') return 0;
  return json_array_not_all_strings(b, (size_t)(e - b)) ? -1 : 0;
}

/* newNodeFromLine: returns -1 on parse error; otherwise sets *kind: 0 other, 1 from, 2 arg */
static int node_from_line(const char* line, size_t n, const directive* d, int* kind) {
  const uint8_t* b = (const uint8_t*)line; const uint8_t* e = b + n;
  trim_space(&b, &e);
  const uint8_t* p = b;
  while (p < e && !is_ascii_ws(*p)) p++;
  const char* cmd = (const char*)b; size_t cn = (size_t)(p - b);
  const uint8_t* rest_b = p; const uint8_t* rest_e = e;
  int has_rest = p < e;
  if (has_rest) {
    while (rest_b < e && is_ascii_ws(*rest_b)) rest_b++;
    size_t off = skip_builder_flags((const char*)rest_b, (size_t)(rest_e - rest_b));
    rest_b += off;
    trim_space(&rest_b, &rest_e);
  } else rest_b = rest_e = e;
  *kind = cmd_is(cmd, cn, "from") ? 1 : cmd_is(cmd, cn, "arg") ? 2 : 0;
  size_t rn = (size_t)(rest_e - rest_b);
  if (cmd_is(cmd, cn, "env") || cmd_is(cmd, cn, "label")) { /* parseNameVal */
    int f, a;
    size_t words = parse_words((const char*)rest_b, rn, d->escape, &f, &a);
    if (words == 0) return 0;
    if (!f) { /* old format: KEY name value -- tokenWhitespace.Split(rest, 2) must yield 2 parts */
      for (size_t i = 0; i < rn; i++) if (is_ascii_ws(rest_b[i])) return 0;
      return -1; /* "<KEY> must have two arguments" */
    }
    return a ? 0 : -1; /* "Syntax error - can't find = in ..." */
  }
  if (cmd_is(cmd, cn, "onbuild")) { /* parseSubCommand */
    if (rn == 0) return 0;
    int k2;
    return node_from_line((const char*)rest_b, rn, d, &k2);
  }
  /* the JSON-form commands: parseMaybeJSON (cmd, entrypoint, run, shell), parseMaybeJSONToList (add, copy, volume) */
  if (cmd_is(cmd, cn, "add") || cmd_is(cmd, cn, "cmd") || cmd_is(cmd, cn, "copy") || cmd_is(cmd, cn, "entrypoint") || cmd_is(cmd, cn, "run") ||
      cmd_is(cmd, cn, "shell") || cmd_is(cmd, cn, "volume"))
    return maybe_json(rest_b, rest_e);
  if (cmd_is(cmd, cn, "healthcheck")) { /* parseHealthConfig: the first word (bytes up to a byte that IsSpace), then parseMaybeJSON */
    const uint8_t* q = rest_b;
    while (q < rest_e && !is_space_rune(*q)) q++;
    if (q == rest_b) return 0;
    while (q < rest_e && is_space_rune(*q)) q++;
    return maybe_json(q, rest_e);
  }
  return 0;
}

/* regexp `(?i)FROM\s+(--platform=[^\s]+)?[^\s]+(\s+AS\s+[^\s]+)?\s*(#.+)?$` (dockerfile2kube.go:131),
 * unanchored at the start: tried at every occurrence of "from". */
static int tail_ok(const uint8_t* s, size_t n, size_t k) { /* \s*(#.+)?$ from k */
  while (k < n && is_re_s(s[k])) k++;
  if (k == n) return 1;
  return s[k] == '#' && k + 1 < n;
}
static int run_has_hash_cut(const uint8_t* s, size_t n, size_t j, size_t k) { /* '#' inside run [j,k) at index >= j+1 with >=1 char after */
  for (size_t p = j + 1; p < k; p++) if (s[p] == '#' && p + 1 < n) return 1;
  return 0;
}
static int from_regex(const uint8_t* s, size_t n) {
  for (size_t i = 0; i + 4 <= n; i++) {
    if (strncasecmp((const char*)s + i, "from", 4) != 0) continue;
    size_t j = i + 4;
    if (j >= n || !is_re_s(s[j])) continue;
    while (j < n && is_re_s(s[j])) j++;
    if (j == n) continue;
    size_t k = j;
    while (k < n && !is_re_s(s[k])) k++;
    if (run_has_hash_cut(s, n, j, k)) return 1;
    if (tail_ok(s, n, k)) return 1;
    /* (\s+AS\s+[^\s]+) ; under (?i) 'S' also matches U+017F */
    size_t a = k;
    while (a < n && is_re_s(s[a])) a++;
    if (a > k && a < n && (s[a] == 'a' || s[a] == 'A')) {
      size_t b = a + 1, sw = 0;
      if (b < n && (s[b] == 's' || s[b] == 'S')) sw = 1;
      else if (b + 1 < n && s[b] == 0xC5 && s[b + 1] == 0xBF) sw = 2;
      if (sw) {
        size_t c = b + sw;
        if (c < n && is_re_s(s[c])) {
          while (c < n && is_re_s(s[c])) c++;
          if (c < n) {
            size_t e2 = c;
            while (e2 < n && !is_re_s(s[e2])) e2++;
            if (run_has_hash_cut(s, n, c, e2)) return 1;
            if (tail_ok(s, n, e2)) return 1;
          }
        }
      }
    }
  }
  return 0;
}

/* Returns M2K_FILE_* bits for the Dockerfile half: FROM_FIRST, DOCKERFILE, LONG_LINE. */
uint32_t m2ko_dockerfile_flags(const uint8_t* data, uint64_t len) {
  directive d = {'\\', 0, 0};
  uint32_t flags = 0;
  int parse_error = 0, n_nodes = 0, first_kind = -1, first_from_match = 0, decided = 0;
  uint64_t pos = 0; int current_line = 0;
  sbuf line = {0, 0, 0}, cont = {0, 0, 0};
  /* scanner: next physical line -> [lb, le); 0 at EOF.  The 64 KiB token cap is checked separately
   * below (any such line fails Parse), so the instruction bits are defined for every file. */
#define NEXT_LINE(lb, le, rc) do { \
    if (pos >= len) { rc = 0; break; } \
    const uint8_t* nl_ = (const uint8_t*)memchr(data + pos, '\n', len - pos); \
    uint64_t end_ = nl_ ? (uint64_t)(nl_ - data) : len; \
    lb = data + pos; le = data + end_; if (le > lb && le[-1] == '\r') le--; \
    pos = end_ + 1; rc = 1; } while (0)
  for (;;) {
    const uint8_t *lb = NULL, *le = NULL; int rc;
    NEXT_LINE(lb, le, rc);
    if (rc == 0) break;
    if (current_line == 0 && le - lb >= 3 && lb[0] == 0xEF && lb[1] == 0xBB && lb[2] == 0xBF) lb += 3;
    /* processLine(stripLeftWhitespace=true) */
    while (lb < le) { size_t w; uint32_t r = decode_rune(lb, (size_t)(le - lb), &w); if (!is_space_rune(r)) break; lb += w; }
    if (possible_directive(&d, (const char*)lb, (size_t)(le - lb)) < 0) { parse_error = 1; break; }
    line.n = 0; if (!line.p) sb_add(&line, "", 0);
    if (!(lb < le && *lb == '#')) sb_add(&line, (const char*)lb, (size_t)(le - lb)); else line.p[0] = 0;
    current_line++;
    int eol = trim_continuation(&d, &line);
    if (eol && line.n == 0) continue;
    while (!eol) {
      NEXT_LINE(lb, le, rc);
      if (rc == 0) break;
      if (possible_directive(&d, (const char*)lb, (size_t)(le - lb)) < 0) { parse_error = 1; break; }
      current_line++;
      /* isComment(scanner.Bytes()) = tokenComment.Match(trimWhitespace(line)): the raw line, left-trimmed first -- an
       * INDENTED comment inside a continuation is dropped too (buildkit's TestParseWarnsOnEmptyContinutationLine has
       * "RUN indented \\ / <4 blanks># this is an indented comment / comment"); then isEmptyContinuationLine */
      { const uint8_t* q = lb; while (q < le) { size_t w; uint32_t r = decode_rune(q, (size_t)(le - q), &w); if (!is_space_rune(r)) break; q += w; }
        if (q == le || *q == '#') continue; }
      cont.n = 0; sb_add(&cont, (const char*)lb, (size_t)(le - lb));
      eol = trim_continuation(&d, &cont);
      sb_add(&line, cont.p, cont.n);
    }
    if (parse_error) break;
    int kind;
    if (node_from_line(line.p, line.n, &d, &kind) < 0) { parse_error = 1; break; }
    n_nodes++;
    if (!decided) {
      if (kind == 2) { /* arg: keep looking */ }
      else { decided = 1; first_kind = kind; if (kind == 1) first_from_match = from_regex((const uint8_t*)line.p, line.n); }
    }
  }
#undef NEXT_LINE
  /* a line of >= 65536 bytes anywhere fails Parse (bufio.ErrTooLong -> handleScannerError) */
  {
    uint64_t p2 = 0;
    while (p2 < len) {
      const uint8_t* nl = (const uint8_t*)memchr(data + p2, '\n', len - p2);
      uint64_t end = nl ? (uint64_t)(nl - data) : len;
      if (end - p2 >= 65536) { flags |= M2K_FILE_LONG_LINE; break; }
      p2 = end + 1;
    }
  }
  free(line.p); free(cont.p);
  /* FROM_FIRST is reported only for files without an over-long line (Parse fails on those anyway) */
  if (decided && first_kind == 1 && first_from_match && !(flags & M2K_FILE_LONG_LINE)) flags |= M2K_FILE_FROM_FIRST;
  if (!parse_error && !(flags & M2K_FILE_LONG_LINE) && n_nodes > 0 && (flags & M2K_FILE_FROM_FIRST)) flags |= M2K_FILE_DOCKERFILE;
  return flags;
}

uint32_t m2ko_file_flags(const uint8_t* data, uint64_t len) {
  uint32_t f = m2ko_dockerfile_flags(data, len);
  if (has_main(data, len)) f |= M2K_FILE_HAS_MAIN;
  return f;
}

/* threaded scan of files [f0, f1): the CPU baseline of the byte-bound half */
typedef struct { const m2k_tree* t; uint32_t f0, f1; uint8_t* out; int tid, nt; volatile uint32_t* next; } scanjob;
static void* scan_worker(void* arg) {
  scanjob* j = (scanjob*)arg;
  for (;;) {
    uint32_t f = __sync_fetch_and_add(j->next, 64);
    if (f >= j->f1) break;
    uint32_t fe = f + 64 < j->f1 ? f + 64 : j->f1;
    for (; f < fe; f++) j->out[f - j->f0] = (uint8_t)m2ko_file_flags(j->t->bytes + j->t->file_off[f], j->t->file_len[f]);
  }
  return NULL;
}
int m2ko_scan_files(const m2k_tree* t, uint32_t f0, uint32_t f1, uint8_t* out, int n_threads) {
  if (!t->bytes || f1 > t->n_files || f0 > f1) return M2K_ERR_ARG;
  if (n_threads < 1) n_threads = 1;
  volatile uint32_t next = f0;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)n_threads);
  scanjob* jobs = (scanjob*)malloc(sizeof(scanjob) * (size_t)n_threads);
  for (int i = 0; i < n_threads; i++) {
    jobs[i] = (scanjob){t, f0, f1, out, i, n_threads, &next};
    pthread_create(&th[i], NULL, scan_worker, &jobs[i]);
  }
  for (int i = 0; i < n_threads; i++) pthread_join(th[i], NULL);
  free(th); free(jobs);
  return M2K_OK;
}

/* ------------------------------------------------------------------------------------------
 * Any2KubeTranslator.GetServiceOptions -- internal/source/any2kube.go:43-105
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const m2k_tree* t; int sh; strlist pre, ign_d, ign_c; m2ko_result* r;
} walkctx;

static int ends_py(const m2k_tree* t, uint32_t e) { /* awk '/.py$/' on the full path: basename ends "py" */
  return name_ends(t, e, "py");
}

static void walk_dir(walkctx* w, uint32_t d) {
  const m2k_tree* t = w->t;
  m2ko_result* r = w->r;
  if (t->dir_flags[d] & 1) { r->dir_state[d] = M2K_DIR_WALK_ERROR; return; } /* :59-62 (ReadDir error reported once) */
  char* path = m2ko_dir_path(t, d);
  int in_pre = is_string_present(&w->pre, path);
  int in_d = is_string_present(&w->ign_d, path);
  int in_c = is_string_present(&w->ign_c, path);
  r->dir_ignore[d] = (uint8_t)((in_d ? 1 : 0) | (in_c ? 2 : 0) | (in_pre ? 4 : 0));
  free(path);
  int descend = 0;
  if (in_pre) r->dir_state[d] = M2K_DIR_PRECONTAINERIZED;                 /* :66-68 */
  else if (in_d) {                                                         /* :69-74 */
    if (in_c) r->dir_state[d] = M2K_DIR_IGNORED_PRUNE;
    else { r->dir_state[d] = M2K_DIR_IGNORED_DESCEND; descend = 1; }
  } else {
    uint32_t opts = m2ko_dir_options(t, d, w->sh);                         /* :75 */
    if (!opts) {
      if (in_c) r->dir_state[d] = M2K_DIR_NOMATCH_PRUNE;                   /* :78-80 */
      else { r->dir_state[d] = M2K_DIR_NOMATCH_DESCEND; descend = 1; }     /* :81 */
    } else {
      r->dir_state[d] = M2K_DIR_SERVICE;                                   /* :83-97 */
      r->service_dirs[r->n_services++] = d;
    }
  }
  if (!descend) return;
  for (uint32_t e = t->dir_first_entry[d]; e < t->dir_first_entry[d + 1]; e++)
    if ((t->entry_type[e] & 0xF) == M2K_T_DIR) walk_dir(w, t->entry_child[e]);
}

static uint32_t main_file_rec(const m2k_tree* t, const uint8_t* fflags, uint32_t d) {
  uint32_t best = M2K_NONE;
  for (uint32_t e = t->dir_first_entry[d]; e < t->dir_first_entry[d + 1]; e++) {
    if ((t->entry_type[e] & 0xF) == M2K_T_DIR) { uint32_t b = main_file_rec(t, fflags, t->entry_child[e]); if (b < best) best = b; }
    else if ((t->entry_type[e] >> 4) == M2K_T_REG && ends_py(t, e)) {
      uint32_t f = t->entry_child[e];
      if ((fflags[f] & M2K_FILE_HAS_MAIN) && f < best) best = f;
    }
  }
  return best;
}

int m2ko_detect(const m2k_tree* t, int sh_flavor, uint32_t n_pre, const char* const* pre_paths, int scan_content,
                int n_threads, m2ko_result** out) {
  m2ko_result* r = (m2ko_result*)calloc(1, sizeof *r);
  r->n_dirs = t->n_dirs; r->n_files = t->n_files;
  r->dir_opts = (uint32_t*)calloc(t->n_dirs ? t->n_dirs : 1, 4);
  r->dir_state = (uint8_t*)calloc(t->n_dirs ? t->n_dirs : 1, 1);
  r->dir_ignore = (uint8_t*)calloc(t->n_dirs ? t->n_dirs : 1, 1);
  r->dir_main_file = (uint32_t*)malloc((t->n_dirs ? t->n_dirs : 1) * 4);
  r->service_dirs = (uint32_t*)calloc(t->n_dirs ? t->n_dirs : 1, 4);
  r->file_flags = (uint8_t*)calloc(t->n_files ? t->n_files : 1, 1);
  walkctx w; memset(&w, 0, sizeof w);
  w.t = t; w.sh = sh_flavor; w.r = r;
  for (uint32_t i = 0; i < n_pre; i++) sl_push(&w.pre, strdup(pre_paths[i]));
  get_ignore_paths(t, &w.ign_d, &w.ign_c);
  if (scan_content && t->bytes && t->n_files) m2ko_scan_files(t, 0, t->n_files, r->file_flags, n_threads);
  for (uint32_t d = 0; d < t->n_dirs; d++) {
    r->dir_opts[d] = m2ko_dir_options(t, d, sh_flavor);
    r->dir_main_file[d] = M2K_NONE;
  }
  if (scan_content) for (uint32_t d = 0; d < t->n_dirs; d++) r->dir_main_file[d] = main_file_rec(t, r->file_flags, d);
  if (t->n_dirs) walk_dir(&w, 0);
  r->n_ignore_dirs = (uint32_t)w.ign_d.n; r->ignore_dirs = w.ign_d.s;
  r->n_ignore_contents = (uint32_t)w.ign_c.n; r->ignore_contents = w.ign_c.s;
  sl_free(&w.pre);
  *out = r;
  return M2K_OK;
}

void m2ko_result_free(m2ko_result* r) {
  if (!r) return;
  free(r->dir_opts); free(r->dir_state); free(r->dir_ignore); free(r->dir_main_file); free(r->service_dirs); free(r->file_flags);
  for (uint32_t i = 0; i < r->n_ignore_dirs; i++) free(r->ignore_dirs[i]);
  for (uint32_t i = 0; i < r->n_ignore_contents; i++) free(r->ignore_contents[i]);
  free(r->ignore_dirs); free(r->ignore_contents);
  free(r);
}
void m2ko_free(void* p) { free(p); }
