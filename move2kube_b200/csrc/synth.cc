// Synthetic source trees (BASELINE.md section 4 / SURVEY.md section 8d): structure on the host,
// content procedural (synth.h).  Deterministic in `seed`; nothing touches a disk.
//   config B: 10k files, ~1 GB, 600 dirs, depth <= 6, 20 .m2kignore files
//   config C: 1M files, ~100 GB, 60k dirs, depth <= 12, 1000 .m2kignore files x 10 lines
//   config D: n_huge multi-GiB files + 1M tiny files
// All names are valid UTF-8 without newlines and there are no symlinks (restriction stated in DESIGN.md).
#include <math.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <thread>

#include "m2k_internal.h"
#include "synth.h"

namespace m2k {

namespace {

struct Rng {
  uint64_t s;
  uint64_t next() { s += 0x9E3779B97F4A7C15ull; return splitmix64(s); }
  uint32_t below(uint32_t n) { return n ? (uint32_t)(next() % n) : 0; }
  double unit() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
  double normal() {  // Box-Muller
    double u1 = unit(), u2 = unit();
    if (u1 < 1e-300) u1 = 1e-300;
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
  }
};

enum Lang { L_NONE, L_NODE, L_GO, L_MAVEN, L_PYTHON, L_PHP, L_GRADLE, L_ANT, L_RUBY, L_DJANGO, L_WAR };

struct GDir { uint32_t parent; uint32_t depth; std::string name; Lang lang; std::vector<uint32_t> child_dirs, files; };
struct GFile { uint32_t dir; std::string name; uint64_t len; uint8_t cls; uint64_t inject; int literal; };

const char* kStems[] = {"src", "pkg", "lib", "cmd", "svc", "api", "web", "core", "util", "test", "docs", "internal", "vendor", "app", "mod", "tools"};
const char* kExtByLang[][3] = {
    {".txt", ".md", ".c"}, {".js", ".json", ".ts"}, {".go", ".go", ".mod"}, {".java", ".xml", ".java"}, {".py", ".py", ".cfg"},
    {".php", ".php", ".html"}, {".java", ".gradle", ".kt"}, {".java", ".xml", ".properties"}, {".rb", ".rb", ".erb"},
    {".py", ".html", ".py"}, {".jsp", ".xml", ".class"}};

Lang pick_lang(Rng& r) {
  uint32_t v = r.below(100);
  if (v < 15) return L_NODE;
  if (v < 30) return L_GO;
  if (v < 45) return L_MAVEN;
  if (v < 60) return L_PYTHON;
  if (v < 75) return L_PHP;
  if (v < 80) return L_GRADLE;
  if (v < 85) return L_ANT;
  if (v < 90) return L_RUBY;
  if (v < 95) return L_DJANGO;
  return L_WAR;
}

}  // namespace

int synth_tree_impl(bool pinned, const m2k_synth_params* p, int host_bytes, m2k_tree** out, std::string* err) {
  *out = nullptr;
  if (!p || p->n_dirs == 0) { *err = "synthetic tree needs at least one directory"; return M2K_ERR_ARG; }
  Rng rng{p->seed * 0x2545F4914F6CDD1Dull + 1};
  uint32_t max_depth = p->max_depth ? p->max_depth : 6;
  std::vector<GDir> dirs(p->n_dirs);
  dirs[0] = {M2K_NONE, 0, "", L_NONE, {}, {}};
  for (uint32_t i = 1; i < p->n_dirs; i++) {
    uint32_t par = rng.below(i);
    for (int tries = 0; tries < 8 && dirs[par].depth >= max_depth; tries++) par = rng.below(i);
    while (dirs[par].depth >= max_depth) par = dirs[par].parent;
    GDir& d = dirs[i];
    d.parent = par; d.depth = dirs[par].depth + 1;
    char buf[64];
    uint32_t odd = rng.below(1000);
    if (odd == 0) snprintf(buf, sizeof buf, "gen%u.go", i);          // a DIRECTORY matching find -name "*.go"
    else if (odd == 1) snprintf(buf, sizeof buf, "tpl%u.php", i);
    else snprintf(buf, sizeof buf, "%s%u", kStems[rng.below(16)], i);
    d.name = buf;
    d.lang = (rng.below(100) < 30) ? pick_lang(rng) : L_NONE;
    dirs[par].child_dirs.push_back(i);
  }
  std::vector<GFile> files;
  files.reserve((size_t)p->n_files + 16);
  auto add_file = [&](uint32_t dir, const std::string& name) {
    files.push_back({dir, name, 0, SC_TEXT, ~0ull, -1});
    dirs[dir].files.push_back((uint32_t)files.size() - 1);
  };
  // marker files of project roots
  for (uint32_t i = 1; i < p->n_dirs && files.size() + 3 < p->n_files; i++) {
    switch (dirs[i].lang) {
      case L_NODE: add_file(i, "package.json"); break;
      case L_GO: add_file(i, "go.mod"); add_file(i, "main.go"); break;
      case L_MAVEN: add_file(i, "pom.xml"); break;
      case L_PYTHON: add_file(i, "requirements.txt"); add_file(i, "main.py"); break;
      case L_PHP: add_file(i, "index.php"); break;
      case L_GRADLE: add_file(i, "build.gradle"); break;
      case L_ANT: add_file(i, "build.xml"); break;
      case L_RUBY: add_file(i, "Gemfile"); break;
      case L_DJANGO: add_file(i, "Pipfile"); add_file(i, "manage.py"); break;
      case L_WAR: add_file(i, "app.war"); break;
      default: break;
    }
  }
  // .m2kignore files: exact-path lines (the reference has no glob semantics, any2kube.go:151-181)
  std::vector<std::string> literals;
  if (p->n_ignore_files && files.size() < p->n_files) {
    // root rule file, like samples/.m2kignore: ignore the root itself and most shallow directories
    // (but not their contents), so that the walk reaches the projects deeper down
    std::vector<std::string> rel(p->n_dirs);
    std::string text = ".\n";
    for (uint32_t i = 1; i < p->n_dirs; i++) {
      rel[i] = dirs[i].parent == 0 ? dirs[i].name : rel[dirs[i].parent] + "/" + dirs[i].name;
      if (dirs[i].depth <= 3 && rng.below(100) < 70) text += rel[i] + "/\n";
    }
    add_file(0, ".m2kignore");
    files.back().literal = (int)literals.size();
    files.back().len = text.size();
    literals.push_back(text);
  }
  {
    uint32_t placed = 0;
    for (uint32_t k = 0; k < p->n_ignore_files && files.size() < p->n_files; k++) {
      uint32_t d = rng.below(p->n_dirs);
      bool dup = false;
      for (uint32_t f : dirs[d].files) if (files[f].name == ".m2kignore") dup = true;
      if (dup) continue;
      std::string text;
      for (int ln = 0; ln < 10; ln++) {
        // target: a directory somewhere below d (walk down a few random steps), or d itself
        uint32_t tdir = d;
        std::string rel;
        for (int step = 0; step < 3 && !dirs[tdir].child_dirs.empty(); step++) {
          tdir = dirs[tdir].child_dirs[rng.below((uint32_t)dirs[tdir].child_dirs.size())];
          if (!rel.empty()) rel += '/';
          rel += dirs[tdir].name;
          if (rng.below(2)) break;
        }
        if (rel.empty()) rel = ".";
        uint32_t form = rng.below(100);
        if (form >= 90 && form < 95) for (char& c : rel) if (c >= 'a' && c <= 'z' && rng.below(2)) c = (char)(c - 32);  // EqualFold
        if (form >= 95) rel = "missing" + std::to_string(rng.below(1000)) + "/" + rel;                             // dangling
        if (form < 40 || (form >= 90 && form < 93)) text += rel + "/\n";
        else if (form < 80 || form >= 93) text += rel + "/*\n";
        else text += ".\n";
      }
      add_file(d, ".m2kignore");
      files.back().literal = (int)literals.size();
      files.back().len = text.size();
      literals.push_back(text);
      placed++;
    }
    (void)placed;
  }
  // the rest: ordinary files, extension by the directory's language
  uint32_t n_marker = (uint32_t)files.size();
  for (uint32_t i = n_marker; i < p->n_files; i++) {
    uint32_t d = rng.below(p->n_dirs);
    Lang lg = dirs[d].lang;
    if (lg == L_NONE && dirs[d].parent != M2K_NONE) lg = dirs[dirs[d].parent].lang;
    char buf[64];
    uint32_t special = rng.below(400);
    if (special == 0) snprintf(buf, sizeof buf, "Dockerfile.%u", i);
    else if (special == 1) snprintf(buf, sizeof buf, "deploy%u.yaml", i);
    else if (special == 2) snprintf(buf, sizeof buf, "values%u.yml", i);
    else snprintf(buf, sizeof buf, "f%u%s", i, kExtByLang[lg][rng.below(3)]);
    add_file(d, buf);
  }
  // sizes
  const uint32_t nf = (uint32_t)files.size();
  {
    std::vector<double> raw(nf, 0.0);
    double sum = 0;
    uint32_t huge_left = p->n_huge, large_left = p->n_large;
    for (uint32_t f = 0; f < nf; f++) {
      if (files[f].literal >= 0) continue;
      if (huge_left && f >= n_marker) { files[f].len = p->huge_bytes; huge_left--; continue; }
      if (large_left && f >= n_marker) { files[f].len = p->large_bytes; large_left--; continue; }
      if (p->tiny_max) { files[f].len = p->tiny_min + rng.below(p->tiny_max - p->tiny_min + 1); continue; }
      double v = exp(log(40000.0) + 1.5 * rng.normal());
      if (v > 8.0 * 1024 * 1024) v = 8.0 * 1024 * 1024;
      raw[f] = v; sum += v;
    }
    if (!p->tiny_max && sum > 0) {
      double scale = (double)p->total_bytes / sum;
      for (uint32_t f = 0; f < nf; f++) if (raw[f] > 0) files[f].len = (uint64_t)(raw[f] * scale);
    }
  }
  // classes and "__main__" injection
  for (uint32_t f = 0; f < nf; f++) {
    GFile& g = files[f];
    if (g.literal >= 0) continue;
    uint64_t h = splitmix64(p->seed ^ (0xA24BAED4963EE407ull * (f + 1)));
    uint32_t c = (uint32_t)(h % 1000);
    bool is_py = g.name.size() >= 2 && g.name.compare(g.name.size() - 2, 2, "py") == 0;
    if (g.name.compare(0, 10, "Dockerfile") == 0 || c < 5) g.cls = SC_DOCKERFILE;
    else if (c < 15) g.cls = SC_BINARY;
    else if (c < 18) g.cls = SC_NONL;
    else if (c < 23) g.cls = (uint8_t)(SC_NEAR_ARGS + (c - 18) % 4);
    else g.cls = SC_TEXT;
    if (g.len > (8ull << 20)) g.cls = SC_TEXT;  // config D's multi-GiB files are plain text
    else if (g.cls == SC_DOCKERFILE || g.cls >= SC_NEAR_ARGS) {
      // Dockerfiles are small in real trees: log-normal around 1.5 KB, at most 64 KiB
      double v = exp(log(1500.0) + 1.0 * rng.normal());
      g.len = (uint64_t)(v < 64 ? 64 : v > 65536 ? 65536 : v);
    }
    uint32_t m = (uint32_t)((h >> 20) % 1000);
    bool inject = is_py ? (g.name == "main.py" ? m < 500 : m < 20) : m < 5;
    if (inject && g.len >= 72 && g.cls != SC_BINARY) {
      uint64_t lo = synth_head(g.cls) ? 64 : 0;  // keep template heads intact
      g.inject = lo + (h >> 32) % (g.len - 8 - lo + 1);
    }
  }
  // lay out: DFS pre-order, children and entries in byte order of their names
  struct Ent { const std::string* name; uint32_t ref; bool is_dir; };
  uint32_t nd = p->n_dirs;
  uint64_t name_bytes = 0;
  for (uint32_t d = 1; d < nd; d++) name_bytes += dirs[d].name.size();
  for (uint32_t f = 0; f < nf; f++) name_bytes += files[f].name.size();
  uint32_t ne = (nd - 1) + nf;
  uint64_t total = 0;
  for (uint32_t f = 0; f < nf; f++) total += (files[f].len + 15) & ~15ull;
  total += 16;
  m2k_tree* t = tree_alloc(nd, ne, nf, name_bytes, total, host_bytes != 0, pinned, "/synth");
  if (!t) { *err = "out of host memory for the synthetic tree"; return M2K_ERR_NOMEM; }
  TreeOwner* o = static_cast<TreeOwner*>(t->owner);
  o->synthetic = true; o->synth_seed = p->seed;
  o->synth_class.resize(nf); o->synth_inject.resize(nf);
  std::vector<uint32_t> new_dir_id(nd, M2K_NONE);
  uint32_t next_dir = 0, next_entry = 0, next_file = 0;
  uint64_t next_name = 0, next_off = 0, content = 0;
  uint32_t maxd = 0;
  struct Frame { uint32_t gdir; };
  // iterative pre-order: process a directory (emit its entries), then push child dirs in reverse name order
  std::vector<uint32_t> stack;
  stack.push_back(0);
  std::vector<uint32_t> pending_entry_of_dir(nd, M2K_NONE);  // entry index of each generated dir (set by its parent)
  while (!stack.empty()) {
    uint32_t g = stack.back();
    stack.pop_back();
    uint32_t id = next_dir++;
    new_dir_id[g] = id;
    t->dir_parent[id] = dirs[g].parent == M2K_NONE ? M2K_NONE : new_dir_id[dirs[g].parent];
    t->dir_depth[id] = dirs[g].depth;
    t->dir_entry[id] = pending_entry_of_dir[g];
    if (t->dir_entry[id] != M2K_NONE) t->entry_child[t->dir_entry[id]] = id;
    if (dirs[g].depth > maxd) maxd = dirs[g].depth;
    std::vector<Ent> ents;
    ents.reserve(dirs[g].child_dirs.size() + dirs[g].files.size());
    for (uint32_t c : dirs[g].child_dirs) ents.push_back({&dirs[c].name, c, true});
    for (uint32_t f : dirs[g].files) ents.push_back({&files[f].name, f, false});
    std::sort(ents.begin(), ents.end(), [](const Ent& a, const Ent& b) { return *a.name < *b.name; });
    t->dir_first_entry[id] = next_entry;
    for (const Ent& en : ents) {
      uint32_t e = next_entry++;
      t->entry_dir[e] = id;
      t->entry_name_off[e] = (uint32_t)next_name;
      memcpy(t->names + next_name, en.name->data(), en.name->size());
      next_name += en.name->size();
      if (en.is_dir) {
        t->entry_type[e] = M2K_T_DIR | (M2K_T_DIR << 4);
        t->entry_child[e] = M2K_NONE;  // patched when the directory is numbered
        pending_entry_of_dir[en.ref] = e;
      } else {
        const GFile& gf = files[en.ref];
        uint32_t fid = next_file++;
        t->entry_type[e] = M2K_T_REG | (M2K_T_REG << 4);
        t->entry_child[e] = fid;
        t->file_entry[fid] = e;
        t->file_off[fid] = next_off;
        t->file_len[fid] = gf.len;
        next_off += (gf.len + 15) & ~15ull;
        content += gf.len;
        o->synth_class[fid] = gf.cls;
        o->synth_inject[fid] = gf.inject;
        if (gf.literal >= 0) { o->literal_file.push_back(fid); o->literal_data.push_back(literals[gf.literal]); }
      }
    }
    // children in byte order => push in reverse so the smallest name is numbered next
    for (size_t k = ents.size(); k-- > 0;) if (ents[k].is_dir) stack.push_back(ents[k].ref);
  }
  t->dir_first_entry[nd] = next_entry;
  t->entry_name_off[ne] = (uint32_t)next_name;
  // subtree ends: ids are pre-order, so end[d] = max over descendants + 1; compute bottom-up
  for (uint32_t d = 0; d < nd; d++) t->dir_end[d] = d + 1;
  for (uint32_t d = nd; d-- > 1;) { uint32_t par = t->dir_parent[d]; if (t->dir_end[d] > t->dir_end[par]) t->dir_end[par] = t->dir_end[d]; }
  t->max_depth = maxd;
  t->content_bytes = content;
  if (host_bytes) {
    unsigned nt = std::thread::hardware_concurrency();
    synth_fill_host(t, 0, nf, t->bytes, nt ? (int)nt : 4);
    memset(t->bytes + total - 16, 0, 16);
  }
  *out = t;
  return M2K_OK;
}

// Host fill of files [file_begin, file_end) into dst (dst[0] is tree offset file_off[file_begin]).
void synth_fill_host(const m2k_tree* tree, uint32_t file_begin, uint32_t file_end, uint8_t* dst, int n_threads) {
  const TreeOwner* o = static_cast<const TreeOwner*>(tree->owner);
  if (file_begin >= file_end) return;
  const uint64_t base = tree->file_off[file_begin];
  std::atomic<uint32_t> next{file_begin};
  // large files are split so that config D's multi-GiB files use every thread
  struct Piece { uint32_t f; uint64_t u0, u1; };
  std::vector<Piece> pieces;
  const uint64_t kPieceUnits = (8u << 20) / 16;
  for (uint32_t f = file_begin; f < file_end; f++) {
    uint64_t units = (tree->file_len[f] + 15) >> 4;
    if (units == 0) continue;
    for (uint64_t u = 0; u < units; u += kPieceUnits) pieces.push_back({f, u, std::min(units, u + kPieceUnits)});
  }
  std::atomic<size_t> pi{0};
  auto worker = [&]() {
    for (;;) {
      size_t k = pi.fetch_add(1);
      if (k >= pieces.size()) break;
      const Piece& pc = pieces[k];
      uint64_t key = synth_file_key(o->synth_seed, pc.f);
      uint8_t* d = dst + (tree->file_off[pc.f] - base);
      for (uint64_t u = pc.u0; u < pc.u1; u++) synth_block16(key, o->synth_class[pc.f], tree->file_len[pc.f], o->synth_inject[pc.f], u, d + u * 16);
    }
  };
  if (n_threads < 1) n_threads = 1;
  std::vector<std::thread> th;
  for (int i = 0; i < n_threads; i++) th.emplace_back(worker);
  for (auto& x : th) x.join();
  for (size_t k = 0; k < o->literal_file.size(); k++) {
    uint32_t f = o->literal_file[k];
    if (f < file_begin || f >= file_end) continue;
    uint8_t* d = dst + (tree->file_off[f] - base);
    uint64_t padded = (tree->file_len[f] + 15) & ~15ull;
    memset(d, 0, padded);
    memcpy(d, o->literal_data[k].data(), o->literal_data[k].size());
  }
  (void)next;
}

}  // namespace m2k
