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
#include <condition_variable>
#include <memory>
#include <mutex>
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

// ---- phase 1: the walk, one task per directory, on a team of threads ---------------------------------
// A directory task opens the directory, reads and sorts its names (byte order == sort.Strings), lstats
// (and, for symlinks, stats) every entry relative to the open directory, and queues one task per real
// sub-directory.  Nothing is numbered yet: pre-order ids need the whole shape of the tree.
struct Ent { std::string name; uint8_t type; uint64_t size; uint32_t child_node; };

struct DirNode {
  std::string path;  // the path filepath.Walk reports
  uint32_t depth = 0;
  uint8_t flags = 0;
  std::vector<Ent> ents;
};

static uint8_t mode_type(mode_t m) {
  return S_ISREG(m) ? M2K_T_REG : S_ISDIR(m) ? M2K_T_DIR : S_ISLNK(m) ? M2K_T_LNK : M2K_T_OTHER;
}

struct ParallelWalk {
  std::mutex m;
  std::condition_variable cv;
  std::vector<std::unique_ptr<DirNode>> nodes;
  std::vector<uint32_t> pending;
  unsigned active = 0;
  std::atomic<bool> newline_name{false};

  uint32_t add_node(std::string path, uint32_t depth) {  // caller holds m
    nodes.emplace_back(new DirNode());
    nodes.back()->path = std::move(path);
    nodes.back()->depth = depth;
    return (uint32_t)nodes.size() - 1;
  }

  void process(uint32_t id) {
    DirNode* n;
    { std::lock_guard<std::mutex> lk(m); n = nodes[id].get(); }
    int fd = open(n->path.c_str(), O_RDONLY | O_DIRECTORY | O_CLOEXEC);
    DIR* dp = fd >= 0 ? fdopendir(fd) : nullptr;
    if (!dp) {
      if (fd >= 0) close(fd);
      n->flags = 1;  // ReadDir failed: the walk reports the directory and moves on
      return;
    }
    std::vector<Ent>& ents = n->ents;
    while (struct dirent* de = readdir(dp)) {
      const char* nm = de->d_name;
      if (nm[0] == '.' && (nm[1] == 0 || (nm[1] == '.' && nm[2] == 0))) continue;
      ents.push_back({nm, 0, 0, M2K_NONE});
    }
    std::sort(ents.begin(), ents.end(), [](const Ent& a, const Ent& b) { return a.name < b.name; });
    std::vector<uint32_t> subdirs;
    for (size_t k = 0; k < ents.size(); k++) {
      Ent& en = ents[k];
      struct stat st, st2;
      uint8_t lt = M2K_T_ERR, ft = M2K_T_ERR;
      if (fstatat(fd, en.name.c_str(), &st, AT_SYMLINK_NOFOLLOW) == 0) {
        lt = mode_type(st.st_mode);
        if (lt == M2K_T_LNK) { if (fstatat(fd, en.name.c_str(), &st2, 0) == 0) { ft = mode_type(st2.st_mode); st = st2; } }
        else ft = lt;
      }
      en.type = (uint8_t)(lt | (ft << 4));
      if (lt != M2K_T_DIR && ft == M2K_T_REG) en.size = (uint64_t)st.st_size;
      if (en.name.find('\n') != std::string::npos) newline_name = true;
      if (lt == M2K_T_DIR) subdirs.push_back((uint32_t)k);
    }
    closedir(dp);
    if (!subdirs.empty()) {
      std::lock_guard<std::mutex> lk(m);
      for (uint32_t k : subdirs) {
        std::string cp = n->path;
        if (cp.empty() || cp.back() != '/') cp += '/';
        cp += ents[k].name;
        ents[k].child_node = add_node(std::move(cp), n->depth + 1);
        pending.push_back(ents[k].child_node);
      }
      cv.notify_all();
    }
  }

  void worker() {
    for (;;) {
      uint32_t id;
      {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return !pending.empty() || active == 0; });
        if (pending.empty()) return;  // nothing queued and nobody working: the walk is complete
        id = pending.back();
        pending.pop_back();
        active++;
      }
      process(id);
      {
        std::lock_guard<std::mutex> lk(m);
        active--;
        if (active == 0 && pending.empty()) cv.notify_all();
      }
    }
  }

  void run(const char* root_path, unsigned nt) {
    { std::lock_guard<std::mutex> lk(m); pending.push_back(add_node(root_path, 0)); }
    std::vector<std::thread> th;
    for (unsigned i = 1; i < nt; i++) th.emplace_back([this] { worker(); });
    worker();
    for (auto& x : th) x.join();
  }
};

// ---- phase 2: number the directories in DFS pre-order and flatten into the struct of arrays --------
struct Flat {
  std::vector<uint32_t> dir_parent, dir_end, dir_entry, dir_depth, dir_node;
  std::vector<uint8_t> dir_flags;
  std::vector<uint32_t> entry_dir, entry_name_off, entry_child;
  std::vector<uint8_t> entry_type;
  std::vector<uint8_t> names;
  std::vector<uint32_t> file_entry;
  std::vector<uint64_t> file_len;
  uint32_t max_depth = 0;
};

static void flatten(ParallelWalk& w, Flat& o) {
  size_t n_ents = 0, n_name = 0;
  for (auto& n : w.nodes) { n_ents += n->ents.size(); for (const Ent& e : n->ents) n_name += e.name.size(); }
  const size_t nd = w.nodes.size();
  o.dir_parent.reserve(nd); o.dir_end.assign(nd, 0); o.dir_entry.reserve(nd); o.dir_depth.reserve(nd); o.dir_node.reserve(nd); o.dir_flags.reserve(nd);
  o.entry_dir.reserve(n_ents); o.entry_name_off.reserve(n_ents + 1); o.entry_child.reserve(n_ents); o.entry_type.reserve(n_ents);
  o.names.reserve(n_name);
  struct Frame { uint32_t node, id, e0, next; };  // next: index of the next entry of this directory to look at for a sub-directory
  std::vector<Frame> st;
  auto enter = [&](uint32_t node, uint32_t parent, uint32_t my_entry) {
    const DirNode& n = *w.nodes[node];
    const uint32_t id = (uint32_t)o.dir_parent.size();
    o.dir_parent.push_back(parent); o.dir_entry.push_back(my_entry); o.dir_depth.push_back(n.depth); o.dir_flags.push_back(n.flags);
    o.dir_node.push_back(node);
    if (n.depth > o.max_depth) o.max_depth = n.depth;
    const uint32_t e0 = (uint32_t)o.entry_dir.size();
    for (const Ent& en : n.ents) {
      const uint32_t e = (uint32_t)o.entry_dir.size();
      o.entry_dir.push_back(id);
      o.entry_name_off.push_back((uint32_t)o.names.size());
      o.names.insert(o.names.end(), en.name.begin(), en.name.end());
      o.entry_type.push_back(en.type);
      uint32_t child = M2K_NONE;
      const uint32_t lt = en.type & 0xF, ft = en.type >> 4;
      if (lt != M2K_T_DIR && ft == M2K_T_REG) {
        child = (uint32_t)o.file_entry.size();
        o.file_entry.push_back(e);
        o.file_len.push_back(en.size);
      }
      o.entry_child.push_back(child);
    }
    st.push_back({node, id, e0, 0});
  };
  enter(0, M2K_NONE, M2K_NONE);
  while (!st.empty()) {
    Frame& f = st.back();
    const DirNode& n = *w.nodes[f.node];
    bool descended = false;
    while (f.next < n.ents.size()) {
      const uint32_t k = f.next++;
      if ((n.ents[k].type & 0xF) != M2K_T_DIR) continue;
      const uint32_t e = f.e0 + k;
      o.entry_child[e] = (uint32_t)o.dir_parent.size();
      const uint32_t my_id = f.id;  // `f` dangles once enter() grows the stack
      enter(n.ents[k].child_node, my_id, e);
      descended = true;
      break;
    }
    if (!descended) { o.dir_end[f.id] = (uint32_t)o.dir_parent.size(); st.pop_back(); }
  }
}

}  // namespace

// Everything of a packed tree except the bytes: the walk, the numbering, the layout of the byte blob.
struct PackPlan::Impl {
  ParallelWalk pw;
  Flat w;
  std::vector<uint64_t> off;
  uint64_t total = 0;
  unsigned nt = 4;
};
PackPlan::PackPlan() : p(new Impl()) {}
PackPlan::~PackPlan() { delete p; }

int pack_plan(const char* root_path, int read_content, bool pinned, bool with_bytes, PackPlan* plan, m2k_tree** out, std::string* err) {
  *out = nullptr;
  // filepath.Walk Lstats its root: a root that is not a real directory (a file, a symlink -- even one
  // to a directory) is handed to the callback once, as itself, and nothing below it is looked at
  // (any2kube.go:58-65, utils.go:93-101).  Such a root has no packed form; the caller answers for it.
  struct stat st;
  if (lstat(root_path, &st) != 0) {
    *err = std::string("cannot stat root ") + root_path + ": " + strerror(errno);
    return M2K_ERR_IO;
  }
  if (!S_ISDIR(st.st_mode)) {
    *err = std::string("root ") + root_path + " is not a directory (filepath.Walk visits it alone, as a non-directory)";
    return M2K_ERR_NOTDIR;
  }
  PackPlan::Impl& P = *plan->p;
  unsigned nt = std::thread::hardware_concurrency();
  if (nt == 0) nt = 4;
  if (nt > 32) nt = 32;
  P.nt = nt;
  ParallelWalk& pw = P.pw;
  pw.run(root_path, nt);
  if (pw.newline_name) {
    *err = "a file name contains a newline: the reference's `find | wc -l` predicates are order dependent for such trees";
    return M2K_ERR_UNSUPPORTED;
  }
  Flat& w = P.w;
  flatten(pw, w);
  uint32_t nd = (uint32_t)w.dir_parent.size(), ne = (uint32_t)w.entry_dir.size(), nf = (uint32_t)w.file_entry.size();
  // sizes can change between lstat and read: the read pass is authoritative, so lay files
  // out with the stat sizes and clamp what is read
  // read_content == 0: names only -- but the .m2kignore files are always read, the walk needs their rules
  if (!read_content) {
    for (uint32_t f = 0; f < nf; f++) {
      uint32_t e = w.file_entry[f];
      uint32_t n0 = w.entry_name_off[e], n1 = e + 1 < ne ? w.entry_name_off[e + 1] : (uint32_t)w.names.size();
      if (!(n1 - n0 == 10 && memcmp(w.names.data() + n0, ".m2kignore", 10) == 0)) w.file_len[f] = 0;
    }
  }
  std::vector<uint64_t>& off = P.off;
  off.resize(nf);
  uint64_t total = 0;
  for (uint32_t f = 0; f < nf; f++) { off[f] = total; total += (w.file_len[f] + 15) & ~15ull; }
  total += 16;
  P.total = total;
  m2k_tree* t = tree_alloc(nd, ne, nf, w.names.size(), total, with_bytes, pinned, root_path);
  if (!t) { *err = "out of host memory while packing"; return M2K_ERR_NOMEM; }
  t->max_depth = w.max_depth;
  memcpy(t->dir_parent, w.dir_parent.data(), 4ull * nd); memcpy(t->dir_end, w.dir_end.data(), 4ull * nd);
  memcpy(t->dir_entry, w.dir_entry.data(), 4ull * nd); memcpy(t->dir_depth, w.dir_depth.data(), 4ull * nd);
  memcpy(t->dir_flags, w.dir_flags.data(), nd);
  {
    // entries were appended directory by directory in id order: the CSR is a count per directory
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
  uint64_t cb = 0;
  for (uint32_t f = 0; f < nf; f++) { t->file_off[f] = off[f]; t->file_len[f] = w.file_len[f]; cb += w.file_len[f]; }
  t->content_bytes = cb;
  *out = t;
  return M2K_OK;
}

uint64_t pack_plan_bytes(const PackPlan* plan) { return plan->p->total; }

// Bytes [lo, hi) of the blob the plan laid out, read from the files into dst (dst[0] is byte lo): this
// thread's share of a block that a team fills together.  Files are opened relative to their directory
// and read with pread, so a file that straddles the range contributes just its part; padding is zeroed.
// Returns false when a file turned out shorter than its lstat size (the plan is then stale).
bool pack_read_range(const PackPlan* plan, uint64_t lo, uint64_t hi, uint8_t* dst) {
  const PackPlan::Impl& P = *plan->p;
  const Flat& w = P.w;
  const uint32_t nf = (uint32_t)w.file_entry.size(), ne = (uint32_t)w.entry_dir.size();
  if (hi > P.total) hi = P.total;
  if (lo >= hi) return true;
  // first file whose padded extent ends after lo
  uint32_t f = (uint32_t)(std::upper_bound(P.off.begin(), P.off.end(), lo) - P.off.begin());
  if (f > 0) f--;
  bool ok = true;
  int dfd = -1;
  uint32_t dfd_dir = M2K_NONE;
  uint64_t at = lo;  // everything in [lo, at) has been written
  for (; f < nf && P.off[f] < hi; f++) {
    const uint64_t fb = P.off[f], len = w.file_len[f], fe = fb + len;
    const uint64_t a = std::max(fb, lo), b = std::min(fe, hi);
    if (a > at) { memset(dst + (at - lo), 0, a - at); at = a; }  // padding before this file's part
    if (b <= a) continue;
    const uint32_t e = w.file_entry[f], d = w.entry_dir[e];
    if (d != dfd_dir) {
      if (dfd >= 0) close(dfd);
      dfd = open(P.pw.nodes[w.dir_node[d]]->path.c_str(), O_RDONLY | O_DIRECTORY | O_CLOEXEC);
      dfd_dir = d;
    }
    const std::string name(reinterpret_cast<const char*>(w.names.data()) + w.entry_name_off[e],
                           (e + 1 < ne ? w.entry_name_off[e + 1] : (uint32_t)w.names.size()) - w.entry_name_off[e]);
    uint64_t got = 0;
    const int fd = dfd >= 0 ? openat(dfd, name.c_str(), O_RDONLY | O_CLOEXEC) : -1;
    if (fd >= 0) {
      while (got < b - a) {
        ssize_t r = pread(fd, dst + (a - lo) + got, b - a - got, (off_t)(a - fb + got));
        if (r <= 0) break;
        got += (uint64_t)r;
      }
      close(fd);
    }
    if (got < b - a) { ok = false; memset(dst + (a - lo) + got, 0, b - a - got); }
    at = b;
  }
  if (at < hi) memset(dst + (at - lo), 0, hi - at);
  if (dfd >= 0) close(dfd);
  return ok;
}

// The rule files' contents, for a tree packed without a host copy of the bytes (m2k_prepare reads them).
void pack_keep_rule_files(const PackPlan* plan, m2k_tree* t) {
  const PackPlan::Impl& P = *plan->p;
  const Flat& w = P.w;
  TreeOwner* o = static_cast<TreeOwner*>(t->owner);
  const uint32_t nf = (uint32_t)w.file_entry.size(), ne = (uint32_t)w.entry_dir.size();
  for (uint32_t f = 0; f < nf; f++) {
    const uint32_t e = w.file_entry[f];
    const uint32_t n0 = w.entry_name_off[e], n1 = e + 1 < ne ? w.entry_name_off[e + 1] : (uint32_t)w.names.size();
    if (!(n1 - n0 == 10 && memcmp(w.names.data() + n0, ".m2kignore", 10) == 0)) continue;
    std::string data((size_t)((w.file_len[f] + 15) & ~15ull), '\0');
    if (!data.empty()) pack_read_range(plan, P.off[f], P.off[f] + data.size(), reinterpret_cast<uint8_t*>(&data[0]));
    data.resize((size_t)w.file_len[f]);
    o->literal_file.push_back(f);  // ascending: host_file_bytes searches this list
    o->literal_data.push_back(std::move(data));
  }
}

int pack_tree_impl(bool pinned, const char* root_path, int read_content, m2k_tree** out, std::string* err) {
  PackPlan plan;
  m2k_tree* t = nullptr;
  if (int rc = pack_plan(root_path, read_content, pinned, true, &plan, &t, err)) return rc;
  PackPlan::Impl& P = *plan.p;
  ParallelWalk& pw = P.pw;
  Flat& w = P.w;
  std::vector<uint64_t>& off = P.off;
  const uint64_t total = P.total;
  unsigned nt = P.nt;
  const uint32_t ne = (uint32_t)w.entry_dir.size(), nf = (uint32_t)w.file_entry.size();
  {
    // ---- phase 3: the bytes.  Work units are runs of consecutive file ids (files of one directory are
    // consecutive, so a worker opens the directory once and the files relative to it) and, for large
    // files, segments read with pread, so that one multi-GB file does not pin the pass to one thread.
    constexpr uint64_t kSegBytes = 32ull << 20;
    constexpr uint32_t kRunFiles = 128;
    struct Unit { uint32_t f0, f1; uint64_t seg_off, seg_len; uint64_t got; };
    std::vector<Unit> units;
    for (uint32_t f = 0; f < nf;) {
      if (w.file_len[f] > kSegBytes) {
        for (uint64_t o = 0; o < w.file_len[f]; o += kSegBytes) units.push_back({f, f + 1, o, std::min(kSegBytes, w.file_len[f] - o), 0});
        f++;
        continue;
      }
      uint32_t g = f;
      uint64_t bytes = 0;
      while (g < nf && g - f < kRunFiles && w.file_len[g] <= kSegBytes && bytes <= kSegBytes) { bytes += w.file_len[g]; g++; }
      units.push_back({f, g, 0, 0, 0});
      f = g;
    }
    if (nt > units.size()) nt = units.empty() ? 1 : (unsigned)units.size();
    std::atomic<size_t> next{0};
    auto name_of = [&](uint32_t e) {
      return std::string(reinterpret_cast<const char*>(w.names.data()) + w.entry_name_off[e],
                         (e + 1 < ne ? w.entry_name_off[e + 1] : (uint32_t)w.names.size()) - w.entry_name_off[e]);
    };
    auto worker = [&]() {
      int dfd = -1;
      uint32_t dfd_dir = M2K_NONE;
      for (;;) {
        size_t u = next.fetch_add(1);
        if (u >= units.size()) break;
        Unit& un = units[u];
        for (uint32_t f = un.f0; f < un.f1; f++) {
          const uint64_t cap = w.file_len[f];
          if (cap == 0) continue;
          const uint32_t e = w.file_entry[f], d = w.entry_dir[e];
          if (d != dfd_dir) {
            if (dfd >= 0) close(dfd);
            dfd = open(pw.nodes[w.dir_node[d]]->path.c_str(), O_RDONLY | O_DIRECTORY | O_CLOEXEC);
            dfd_dir = d;
          }
          uint8_t* dst = t->bytes + off[f];
          const uint64_t lo = un.seg_len ? un.seg_off : 0, want = un.seg_len ? un.seg_len : cap;
          uint64_t got = 0;
          int fd = dfd >= 0 ? openat(dfd, name_of(e).c_str(), O_RDONLY | O_CLOEXEC) : -1;
          if (fd >= 0) {
            while (got < want) {
              ssize_t r = pread(fd, dst + lo + got, want - got, (off_t)(lo + got));
              if (r <= 0) break;
              got += (uint64_t)r;
            }
            close(fd);
          }
          if (un.seg_len) { un.got = got; continue; }  // stitched below
          t->file_len[f] = got;  // unreadable or shrunk: what could be read (0 = no content signatures)
          memset(dst + got, 0, ((cap + 15) & ~15ull) - got);
        }
      }
      if (dfd >= 0) close(dfd);
    };
    std::vector<std::thread> th;
    for (unsigned i = 1; i < nt; i++) th.emplace_back(worker);
    worker();
    for (auto& x : th) x.join();
    // segmented files: the content is the run of complete segments plus the first short one
    for (size_t u = 0; u < units.size();) {
      if (!units[u].seg_len) { u++; continue; }
      const uint32_t f = units[u].f0;
      uint64_t len = 0;
      bool open_run = true;
      for (; u < units.size() && units[u].seg_len && units[u].f0 == f; u++) {
        if (!open_run) continue;
        len += units[u].got;
        if (units[u].got < units[u].seg_len) open_run = false;
      }
      t->file_len[f] = len;
      memset(t->bytes + off[f] + len, 0, ((w.file_len[f] + 15) & ~15ull) - len);
    }
    memset(t->bytes + total - 16, 0, 16);
  }
  uint64_t cb = 0;
  for (uint32_t f = 0; f < nf; f++) cb += t->file_len[f];
  t->content_bytes = cb;
  *out = t;
  return M2K_OK;
}

}  // namespace m2k
