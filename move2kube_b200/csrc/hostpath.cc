// Host-side path and .m2kignore handling with Go semantics.
// Reference: internal/source/any2kube.go:151-181 (getIgnorePaths), internal/common/utils.go:308-315
// (IsStringPresent / strings.EqualFold).  [external] path/filepath.Clean/Join/Dir, strings.TrimSpace and
// bufio.Scanner follow the Go 1.15 standard library's documented behaviour.
#include <string.h>

#include <algorithm>

#include "casefold_table.h"
#include "gostr.h"
#include "m2k_internal.h"

namespace m2k {

// filepath.Clean for Unix paths: lexical only.  Writes into out (capacity >= n + 2), returns the length.
size_t path_clean_buf(const char* p, size_t n, char* out) {
  if (n == 0) { out[0] = '.'; return 1; }
  const bool rooted = p[0] == '/';
  size_t w = 0;
  if (rooted) out[w++] = '/';
  const size_t floor0 = rooted ? 1 : 0;
  size_t dotdot = floor0;  // output below this index is a run of leading ".." components: never popped
  size_t i = floor0;
  while (i < n) {
    if (p[i] == '/') { i++; continue; }
    size_t j = i;
    while (j < n && p[j] != '/') j++;
    const size_t len = j - i;
    if (len == 1 && p[i] == '.') {
      // skip
    } else if (len == 2 && p[i] == '.' && p[i + 1] == '.') {
      if (w > dotdot) {  // drop the last component
        size_t k = w - 1;
        while (k > dotdot && out[k] != '/') k--;
        w = k;
      } else if (!rooted) {
        if (w) out[w++] = '/';
        out[w++] = '.'; out[w++] = '.';
        dotdot = w;
      }
    } else {
      if (w > floor0) out[w++] = '/';
      memcpy(out + w, p + i, len);
      w += len;
    }
    i = j;
  }
  if (w == 0) out[w++] = '.';
  return w;
}

std::string path_clean(const std::string& p) {
  std::string out(p.size() + 2, '\0');
  out.resize(path_clean_buf(p.data(), p.size(), &out[0]));
  return out;
}

// filepath.Join of two elements: empty elements are dropped, the result is cleaned.
std::string path_join(const std::string& a, const std::string& b) {
  if (a.empty() && b.empty()) return "";
  if (a.empty()) return path_clean(b);
  if (b.empty()) return path_clean(a);
  return path_clean(a + "/" + b);
}

// filepath.Dir
std::string path_dir(const std::string& p) {
  size_t s = p.rfind('/');
  return path_clean(s == std::string::npos ? std::string() : p.substr(0, s + 1));
}

bool equal_fold(const std::string& a, const std::string& b) {
  const uint8_t* pa = reinterpret_cast<const uint8_t*>(a.data());
  const uint8_t* pb = reinterpret_cast<const uint8_t*>(b.data());
  size_t i = 0, j = 0;
  while (i < a.size() && j < b.size()) {
    uint32_t wa, wb;
    uint32_t ra = fold_rune(decode_rune(pa + i, a.size() - i, &wa), m2k_casefold_from, m2k_casefold_to, M2K_CASEFOLD_N);
    uint32_t rb = fold_rune(decode_rune(pb + j, b.size() - j, &wb), m2k_casefold_from, m2k_casefold_to, M2K_CASEFOLD_N);
    if (ra != rb) return false;
    i += wa;
    j += wb;
  }
  return i == a.size() && j == b.size();
}

uint64_t hash_fold(uint64_t h, const std::string& s) {
  return hash_fold_bytes(h, reinterpret_cast<const uint8_t*>(s.data()), s.size(), m2k_casefold_from, m2k_casefold_to, M2K_CASEFOLD_N);
}

void fold_runes(const std::string& s, std::vector<uint32_t>* out) {
  out->clear();
  const uint8_t* p = reinterpret_cast<const uint8_t*>(s.data());
  for (size_t i = 0; i < s.size();) {
    uint32_t w;
    out->push_back(fold_rune(decode_rune(p + i, s.size() - i, &w), m2k_casefold_from, m2k_casefold_to, M2K_CASEFOLD_N));
    i += w;
  }
}

// strings.TrimSpace on [b, e)
static void trim_space(const uint8_t*& b, const uint8_t*& e) {
  while (b < e) {
    uint32_t w;
    if (!is_space_rune(decode_rune(b, (size_t)(e - b), &w))) break;
    b += w;
  }
  while (e > b) {
    // utf8.DecodeLastRune: back up over at most three continuation bytes
    const uint8_t* s = e - 1;
    uint32_t r, w = 1;
    if (*s < 0x80) {
      r = *s;
    } else {
      const uint8_t* lim = (e - b > 4) ? e - 4 : b;
      while (s > lim && (*s & 0xC0) == 0x80) s--;
      uint32_t ww;
      r = decode_rune(s, (size_t)(e - s), &ww);
      if (s + ww != e) { r = 0xFFFD; w = 1; } else w = ww;
    }
    if (!is_space_rune(r)) break;
    e -= w;
  }
}

// One .m2kignore file (any2kube.go:158-177): bufio.Scanner lines (64 KiB cap: a longer line ends
// the scan silently), TrimSpace, a trailing '*' (one) marks "ignore contents", Join with the
// file's directory.
void parse_ignore_file(const std::string& file_path, const uint8_t* data, uint64_t len, std::vector<std::string>* ignore_dirs,
                       std::vector<std::string>* ignore_contents) {
  PathArena a;
  parse_ignore_file_arena(path_dir(file_path), data, len, &a);
  for (size_t i = 0; i < a.size(); i++) {
    std::string sp(a.chars.data() + a.off[i], a.off[i + 1] - a.off[i]);
    (a.flag[i] == 2 ? ignore_contents : ignore_dirs)->push_back(sp);
  }
}

// Same, appending the cleaned paths to an arena (no allocation per line).  `dir` is the cleaned
// directory of the file (filepath.Dir of its Walk path); flag 1 = ignoreDirectories, 2 = ignoreContents.
void parse_ignore_file_arena(const std::string& dir, const uint8_t* data, uint64_t len, PathArena* out) {
  if (out->off.empty()) out->off.push_back(0);
  std::string tmp;
  uint64_t pos = 0;
  while (pos < len) {
    const uint8_t* nl = static_cast<const uint8_t*>(memchr(data + pos, '\n', len - pos));
    uint64_t end = nl ? (uint64_t)(nl - data) : len;
    if (end - pos >= kLongLine) break;
    const uint8_t* b = data + pos;
    const uint8_t* e = data + end;
    if (e > b && e[-1] == '\r') e--;
    trim_space(b, e);
    uint8_t flag = 1;
    if (e > b && e[-1] == '*') { e--; flag = 2; }
    // filepath.Join(dir, line): empty elements are dropped, then Clean
    tmp.assign(dir);
    if (e > b) {
      if (!tmp.empty()) tmp += '/';
      tmp.append(reinterpret_cast<const char*>(b), (size_t)(e - b));
    }
    size_t at = out->chars.size();
    out->chars.resize(at + tmp.size() + 2);
    size_t n = tmp.empty() ? 0 : path_clean_buf(tmp.data(), tmp.size(), out->chars.data() + at);
    out->chars.resize(at + n);
    out->off.push_back((uint32_t)out->chars.size());
    out->flag.push_back(flag);
    pos = end + 1;
  }
}

static std::string entry_name(const m2k_tree* t, uint32_t e) {
  return std::string(reinterpret_cast<const char*>(t->names + t->entry_name_off[e]), t->entry_name_off[e + 1] - t->entry_name_off[e]);
}

// The path filepath.Walk reports for directory d: the root exactly as given, below it Join(parent, name).
std::string tree_dir_path(const m2k_tree* t, uint32_t d) {
  if (d == 0) return t->root_path ? t->root_path : "";
  std::vector<uint32_t> chain;
  for (uint32_t x = d; x != 0; x = t->dir_parent[x]) chain.push_back(x);
  std::string p = t->root_path ? t->root_path : "";
  for (size_t k = chain.size(); k-- > 0;) p = path_join(p, entry_name(t, t->dir_entry[chain[k]]));
  return p;
}

// Same string as tree_dir_path for d != 0, without cleaning at every level: entry names contain no
// '/' and are never "." or "..", so appending "/name" to the cleaned root is already clean.
std::string fast_dir_path(const m2k_tree* t, uint32_t d, const std::string& clean_root) {
  uint32_t chain[4096];
  uint32_t n = 0;
  size_t len = clean_root.size() + 1;
  std::vector<uint32_t> big;
  for (uint32_t x = d; x != 0; x = t->dir_parent[x]) {
    uint32_t e = t->dir_entry[x];
    len += t->entry_name_off[e + 1] - t->entry_name_off[e] + 1;
    if (n < 4096) chain[n++] = e; else big.push_back(e);
  }
  std::string p;
  p.reserve(len);
  if (clean_root != ".") p = clean_root;
  auto app = [&](uint32_t e) {
    if (!p.empty() && p.back() != '/') p += '/';
    p.append(reinterpret_cast<const char*>(t->names + t->entry_name_off[e]), t->entry_name_off[e + 1] - t->entry_name_off[e]);
  };
  for (size_t k = big.size(); k-- > 0;) app(big[k]);
  for (uint32_t k = n; k-- > 0;) app(chain[k]);
  return p;
}

// filepath.Walk order over all entries: a directory's entries in name order, descending into each
// sub-directory right after its own entry.
// Puts a set of entry ids into filepath.Walk order without walking the tree.  Directories are numbered
// in Walk order, so an entry's position is fixed by the NEXT directory the walk enters after it:
// for a file that is the first real sub-directory listed after it in its own directory, or, when
// there is none, whatever follows its directory's subtree (dir_end).  Files that share that next
// directory are visited deepest directory first (the walk is unwinding), then by name; a directory
// entry comes after all of them.  Cost: one pass over the matched directories' listings + a sort.
void sort_walk_order(const m2k_tree* t, std::vector<uint32_t>* ents) {
  std::vector<uint32_t>& v = *ents;
  std::sort(v.begin(), v.end());
  v.erase(std::unique(v.begin(), v.end()), v.end());
  struct Key { uint32_t next, unwind, e; };
  std::vector<Key> keys(v.size());
  auto real_dir = [&](uint32_t e) { return (t->entry_type[e] & 0xF) == M2K_T_DIR && t->entry_child[e] != M2K_NONE; };
  uint32_t cur_dir = M2K_NONE, limit = 0, cur_next = 0;
  for (size_t k = v.size(); k-- > 0;) {  // entries are sorted by (directory, name): go backwards inside each directory
    const uint32_t e = v[k], d = t->entry_dir[e];
    if (d != cur_dir) { cur_dir = d; limit = t->dir_first_entry[d + 1]; cur_next = t->dir_end[d]; }
    if (real_dir(e)) {
      cur_next = t->entry_child[e];
      keys[k] = {cur_next, 0xFFFFFFFFu, e};
    } else {
      for (uint32_t i = e + 1; i < limit; i++)
        if (real_dir(i)) { cur_next = t->entry_child[i]; break; }
      keys[k] = {cur_next, t->n_dirs - 1 - d, e};
    }
    limit = e;
  }
  std::sort(keys.begin(), keys.end(), [](const Key& a, const Key& b) {
    if (a.next != b.next) return a.next < b.next;
    if (a.unwind != b.unwind) return a.unwind < b.unwind;
    return a.e < b.e;
  });
  for (size_t k = 0; k < v.size(); k++) v[k] = keys[k].e;
}

}  // namespace m2k
