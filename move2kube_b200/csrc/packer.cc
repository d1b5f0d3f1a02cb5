// Host packer: ONE walk of the input tree into the struct-of-arrays layout of include/m2kdetect.h.
// It stands in for every filepath.Walk the reference does over the same tree during planning
// (utils.go:85-120 three times per Any2Kube call, any2kube.go:58, dockerfile2kube.go:154, and the
// GetFilesByExt walks of SURVEY.md section 8 f1).  Walk semantics kept [external: path/filepath]:
// Lstat (symlinks are never descended), children in byte order, unreadable directories recorded.
#include <dirent.h>
#include <errno.h>
#include <fcntl.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <thread>

#include "m2k_internal.h"

namespace m2k {

static void* host_alloc(size_t bytes, bool pinned) {
  if (bytes == 0) bytes = 16;
  void* p = nullptr;
  if (pinned) {
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  } else {
    p = malloc(bytes);
  }
  return p;
}
static void host_free(void* p, bool pinned) {
  if (!p) return;
  if (pinned) cudaFreeHost(p); else free(p);
}

m2k_tree* tree_alloc(uint32_t n_dirs, uint32_t n_entries, uint32_t n_files, uint64_t n_name_bytes, uint64_t n_bytes,
                     bool with_bytes, bool pinned, const char* root) {
  m2k_tree* t = static_cast<m2k_tree*>(calloc(1, sizeof(m2k_tree)));
  TreeOwner* o = new TreeOwner();
  o->pinned = pinned;
  o->root = root ? root : "";
  t->owner = o;
  t->root_path = o->root.c_str();
  t->n_dirs = n_dirs; t->n_entries = n_entries; t->n_files = n_files;
  t->n_name_bytes = n_name_bytes; t->n_bytes = n_bytes;
  bool ok = true;
  auto A = [&](size_t bytes) { void* p = host_alloc(bytes, pinned); if (!p) ok = false; else memset(p, 0, bytes ? bytes : 16); return p; };
  t->dir_parent = (uint32_t*)A(4ull * n_dirs); t->dir_end = (uint32_t*)A(4ull * n_dirs); t->dir_entry = (uint32_t*)A(4ull * n_dirs);
  t->dir_first_entry = (uint32_t*)A(4ull * (n_dirs + 1)); t->dir_depth = (uint32_t*)A(4ull * n_dirs); t->dir_flags = (uint8_t*)A(n_dirs);
  t->entry_dir = (uint32_t*)A(4ull * n_entries); t->entry_name_off = (uint32_t*)A(4ull * (n_entries + 1));
  t->entry_type = (uint8_t*)A(n_entries); t->entry_child = (uint32_t*)A(4ull * n_entries); t->names = (uint8_t*)A(n_name_bytes);
  t->file_entry = (uint32_t*)A(4ull * n_files); t->file_off = (uint64_t*)A(8ull * n_files); t->file_len = (uint64_t*)A(8ull * n_files);
  if (with_bytes) {
    t->bytes = (uint8_t*)host_alloc(n_bytes, pinned);  // not cleared here: the filler writes every byte
    if (!t->bytes) ok = false;
  }
  if (!ok) { tree_release(t); return nullptr; }
  return t;
}

void tree_release(m2k_tree* t) {
  if (!t) return;
  TreeOwner* o = static_cast<TreeOwner*>(t->owner);
  bool pinned = o && o->pinned;
  void* ptrs[] = {t->dir_parent, t->dir_end, t->dir_entry, t->dir_first_entry, t->dir_depth, t->dir_flags, t->entry_dir,
                  t->entry_name_off, t->entry_type, t->entry_child, t->names, t->file_entry, t->file_off, t->file_len, t->bytes};
  for (void* p : ptrs) host_free(p, pinned);
  delete o;
  free(t);
}

namespace {

struct Ent { std::string name; uint8_t type; };

struct Walker {
  std::vector<uint32_t> dir_parent, dir_end, dir_entry, dir_first, dir_depth;
  std::vector<uint8_t> dir_flags;
  std::vector<std::string> dir_path;
  std::vector<uint32_t> entry_dir, entry_name_off, entry_child;
  std::vector<uint8_t> entry_type;
  std::vector<uint8_t> names;
  std::vector<uint32_t> file_entry;
  std::vector<uint64_t> file_len;
  uint32_t max_depth = 0;
  bool newline_name = false;

  static uint8_t mode_type(mode_t m) {
    return S_ISREG(m) ? M2K_T_REG : S_ISDIR(m) ? M2K_T_DIR : S_ISLNK(m) ? M2K_T_LNK : M2K_T_OTHER;
  }

  void visit(int parent_fd, const char* name, const std::string& path, uint32_t parent, uint32_t my_entry, uint32_t depth) {
    uint32_t id = (uint32_t)dir_parent.size();
    dir_parent.push_back(parent); dir_end.push_back(0); dir_entry.push_back(my_entry); dir_depth.push_back(depth);
    dir_flags.push_back(0); dir_path.push_back(path);
    dir_first.push_back((uint32_t)entry_dir.size());
    if (depth > max_depth) max_depth = depth;
    int fd = openat(parent_fd, name, O_RDONLY | O_DIRECTORY | O_CLOEXEC);
    DIR* dp = fd >= 0 ? fdopendir(fd) : nullptr;
    if (!dp) {
      if (fd >= 0) close(fd);
      dir_flags[id] = 1;  // ReadDir failed: the walk reports the directory and moves on
      dir_end[id] = (uint32_t)dir_parent.size();
      return;
    }
    std::vector<Ent> ents;
    while (struct dirent* de = readdir(dp)) {
      const char* n = de->d_name;
      if (n[0] == '.' && (n[1] == 0 || (n[1] == '.' && n[2] == 0))) continue;
      ents.push_back({n, 0});
    }
    std::sort(ents.begin(), ents.end(), [](const Ent& a, const Ent& b) { return a.name < b.name; });  // byte order == sort.Strings
    uint32_t e0 = (uint32_t)entry_dir.size();
    for (Ent& en : ents) {
      struct stat st, st2;
      uint8_t lt = M2K_T_ERR, ft = M2K_T_ERR;
      if (fstatat(fd, en.name.c_str(), &st, AT_SYMLINK_NOFOLLOW) == 0) {
        lt = mode_type(st.st_mode);
        if (lt == M2K_T_LNK) { if (fstatat(fd, en.name.c_str(), &st2, 0) == 0) { ft = mode_type(st2.st_mode); st = st2; } }
        else ft = lt;
      }
      en.type = (uint8_t)(lt | (ft << 4));
      if (en.name.find('\n') != std::string::npos) newline_name = true;
      uint32_t e = (uint32_t)entry_dir.size();
      entry_dir.push_back(id);
      entry_name_off.push_back((uint32_t)names.size());
      names.insert(names.end(), en.name.begin(), en.name.end());
      entry_type.push_back(en.type);
      uint32_t child = M2K_NONE;
      if (lt != M2K_T_DIR && ft == M2K_T_REG) {
        child = (uint32_t)file_entry.size();
        file_entry.push_back(e);
        file_len.push_back((uint64_t)st.st_size);
      }
      entry_child.push_back(child);
    }
    for (size_t k = 0; k < ents.size(); k++) {
      if ((ents[k].type & 0xF) != M2K_T_DIR) continue;
      uint32_t e = e0 + (uint32_t)k;
      entry_child[e] = (uint32_t)dir_parent.size();
      std::string child_path = path;
      if (child_path.empty() || child_path.back() != '/') child_path += '/';
      child_path += ents[k].name;
      visit(fd, ents[k].name.c_str(), child_path, id, e, depth + 1);
    }
    closedir(dp);
    dir_end[id] = (uint32_t)dir_parent.size();
  }
};

}  // namespace

int pack_tree_impl(bool pinned, const char* root_path, int read_content, m2k_tree** out, std::string* err) {
  *out = nullptr;
  struct stat st;
  if (stat(root_path, &st) != 0) {
    *err = std::string("cannot stat root ") + root_path + ": " + strerror(errno);
    return M2K_ERR_IO;
  }
  Walker w;
  w.visit(AT_FDCWD, root_path, root_path, M2K_NONE, M2K_NONE, 0);
  if (w.newline_name) {
    *err = "a file name contains a newline: the reference's `find | wc -l` predicates are order dependent for such trees";
    return M2K_ERR_UNSUPPORTED;
  }
  uint32_t nd = (uint32_t)w.dir_parent.size(), ne = (uint32_t)w.entry_dir.size(), nf = (uint32_t)w.file_entry.size();
  // sizes can change between lstat and read: the read pass below is authoritative, so lay files
  // out with the stat sizes and clamp what is read
  // read_content == 0: names only -- but the .m2kignore files are always read, the walk needs their rules
  if (!read_content) {
    for (uint32_t f = 0; f < nf; f++) {
      uint32_t e = w.file_entry[f];
      uint32_t n0 = w.entry_name_off[e], n1 = e + 1 < ne ? w.entry_name_off[e + 1] : (uint32_t)w.names.size();
      if (!(n1 - n0 == 10 && memcmp(w.names.data() + n0, ".m2kignore", 10) == 0)) w.file_len[f] = 0;
    }
  }
  std::vector<uint64_t> off(nf);
  uint64_t total = 0;
  for (uint32_t f = 0; f < nf; f++) { off[f] = total; total += (w.file_len[f] + 15) & ~15ull; }
  total += 16;
  m2k_tree* t = tree_alloc(nd, ne, nf, w.names.size(), total, true, pinned, root_path);
  if (!t) { *err = "out of host memory while packing"; return M2K_ERR_NOMEM; }
  t->max_depth = w.max_depth;
  memcpy(t->dir_parent, w.dir_parent.data(), 4ull * nd); memcpy(t->dir_end, w.dir_end.data(), 4ull * nd);
  memcpy(t->dir_entry, w.dir_entry.data(), 4ull * nd); memcpy(t->dir_depth, w.dir_depth.data(), 4ull * nd);
  memcpy(t->dir_flags, w.dir_flags.data(), nd);
  for (uint32_t d = 0; d < nd; d++) t->dir_first_entry[d] = w.dir_first[d];
  t->dir_first_entry[nd] = ne;
  // a directory's entries are contiguous only if nothing was appended between them: true because
  // visit() appends all entries of a directory before recursing; but dir_first of the NEXT directory
  // id is not dir_first[d+1] of the array for non-leaf d.  Recompute the CSR end from entry_dir.
  {
    std::vector<uint32_t> cnt(nd + 1, 0);
    for (uint32_t e = 0; e < ne; e++) cnt[w.entry_dir[e] + 1]++;
    for (uint32_t d = 0; d < nd; d++) cnt[d + 1] += cnt[d];
    for (uint32_t d = 0; d <= nd; d++) t->dir_first_entry[d] = cnt[d];
  }
  memcpy(t->entry_dir, w.entry_dir.data(), 4ull * ne);
  memcpy(t->entry_name_off, w.entry_name_off.data(), 4ull * ne);
  t->entry_name_off[ne] = (uint32_t)w.names.size();
  memcpy(t->entry_type, w.entry_type.data(), ne); memcpy(t->entry_child, w.entry_child.data(), 4ull * ne);
  if (!w.names.empty()) memcpy(t->names, w.names.data(), w.names.size());
  memcpy(t->file_entry, w.file_entry.data(), 4ull * nf);
  for (uint32_t f = 0; f < nf; f++) { t->file_off[f] = off[f]; t->file_len[f] = w.file_len[f]; }
  {
    unsigned nt = std::thread::hardware_concurrency();
    if (nt == 0) nt = 4;
    if (nt > 32) nt = 32;
    if (nt > nf) nt = nf ? nf : 1;
    std::atomic<uint32_t> next{0};
    auto worker = [&]() {
      for (;;) {
        uint32_t f = next.fetch_add(1);
        if (f >= nf) break;
        uint32_t e = w.file_entry[f];
        std::string p = w.dir_path[w.entry_dir[e]];
        if (p.empty() || p.back() != '/') p += '/';
        p.append(reinterpret_cast<const char*>(w.names.data()) + w.entry_name_off[e],
                 (e + 1 < ne ? w.entry_name_off[e + 1] : (uint32_t)w.names.size()) - w.entry_name_off[e]);
        uint64_t cap = w.file_len[f], got = 0;
        uint8_t* dst = t->bytes + off[f];
        if (cap == 0) continue;
        int fd = open(p.c_str(), O_RDONLY | O_CLOEXEC);
        if (fd >= 0) {
          while (got < cap) {
            ssize_t r = read(fd, dst + got, cap - got);
            if (r <= 0) break;
            got += (uint64_t)r;
          }
          close(fd);
        }
        t->file_len[f] = got;  // unreadable or shrunk: what could be read (0 = no content signatures)
        uint64_t padded = (cap + 15) & ~15ull;
        memset(dst + got, 0, padded - got);
      }
    };
    std::vector<std::thread> th;
    for (unsigned i = 0; i < nt; i++) th.emplace_back(worker);
    for (auto& x : th) x.join();
    memset(t->bytes + total - 16, 0, 16);
  }
  uint64_t cb = 0;
  for (uint32_t f = 0; f < nf; f++) cb += t->file_len[f];
  t->content_bytes = cb;
  *out = t;
  return M2K_OK;
}

}  // namespace m2k
