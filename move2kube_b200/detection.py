"""One GPU detection pass over an input tree, cached per process.

The reference creates a fresh `Containerizers` per call and re-walks the tree each time
(any2kube.go:45-46, cfmanifest2kube.go:53-54); here the tree is packed once and every question the
callers ask (options of a directory, files by name / extension, the walk's decisions) is answered
from the arrays one m2k_detect call returned.
"""
import os

from . import _abi, lib

_ctx = None
_cache = {}


def context(device=None):
    """Process-wide m2k context, created lazily (the GPU context must outlive
    the per-call Containerizers objects).  Raises lib.M2KError when there is no CUDA device."""
    global _ctx
    if _ctx is None:
        if device is None:
            device = int(os.environ.get("M2K_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        _ctx = lib.Context(device)
    return _ctx


def reset():
    global _ctx
    for d in list(_cache.values()):
        d.close()
    _cache.clear()
    if _ctx is not None:
        _ctx.close()
        _ctx = None


def sh_flavor():
    """/bin/sh flavour the python detect scripts are evaluated under: bash-as-sh is
    canonical (the project's UBI8 image); M2K_SH=dash selects dash semantics."""
    return _abi.SH_DASH if os.environ.get("M2K_SH", "bash") == "dash" else _abi.SH_BASH


class Detection:
    """Arrays of one m2k_detect call plus path look-ups."""

    def __init__(self, root: str, pre=(), flavor=None, skip_content=False):
        self.root = root
        self.broot = os.fsencode(root)
        self.ctx = context()
        self.flavor = sh_flavor() if flavor is None else flavor
        # one-shot use: m2k_pack_detect walks the tree once and streams the file bytes from the files to the
        # device through the context's pinned blocks -- they are never gathered on the host
        self.tree = None
        self.res = None
        self.tree, self.res = self.ctx.pack_detect(self.broot, self.flavor, [os.fsencode(p) for p in pre], skip_content)
        t, r = self.tree.contents, self.res.contents
        self.n_dirs, self.n_entries, self.n_files = t.n_dirs, t.n_entries, t.n_files
        names = bytes(bytearray(t.names[0:t.n_name_bytes])) if t.n_name_bytes else b""
        offs = t.entry_name_off[0:t.n_entries + 1] if t.n_entries else [0]
        self.entry_name = [names[offs[i]:offs[i + 1]] for i in range(t.n_entries)]
        self.entry_dir = t.entry_dir[0:t.n_entries]
        self.dir_parent = t.dir_parent[0:t.n_dirs]
        self.dir_end = t.dir_end[0:t.n_dirs]
        self.dir_path = [b""] * t.n_dirs
        clean_root = lib.path_clean(self.broot)
        for d in range(t.n_dirs):
            if d == 0:
                self.dir_path[0] = self.broot  # filepath.Walk reports the root as given
                continue
            par = self.dir_parent[d]
            base = clean_root if par == 0 else self.dir_path[par]
            nm = self.entry_name[t.dir_entry[d]]
            if base == b".":
                self.dir_path[d] = nm
            elif base.endswith(b"/"):
                self.dir_path[d] = base + nm
            else:
                self.dir_path[d] = base + b"/" + nm
        self.path_to_dir = {}
        for d in range(t.n_dirs):
            self.path_to_dir.setdefault(self.dir_path[d], d)
        self.dir_opts = r.dir_opts[0:t.n_dirs]
        self.dir_state = bytes(bytearray(r.dir_state[0:t.n_dirs]))
        self.dir_ignore = bytes(bytearray(r.dir_ignore[0:t.n_dirs]))
        self.dir_flags = bytes(bytearray(t.dir_flags[0:t.n_dirs]))
        self.dir_main_file = r.dir_main_file[0:t.n_dirs]
        self.file_flags = bytes(bytearray(r.file_flags[0:t.n_files])) if t.n_files else b""
        self.file_entry = t.file_entry[0:t.n_files]
        self.service_dirs = r.service_dirs[0:r.n_services]
        self.ignore_dirs = [r.ignore_dirs[i] for i in range(r.n_ignore_dirs)]
        self.ignore_contents = [r.ignore_contents[i] for i in range(r.n_ignore_contents)]
        self.timing = {k: getattr(r, k) for k in ("ms_h2d", "ms_kernels", "ms_scan_kernel", "ms_d2h", "ms_host_ignore", "ms_total")}
        self.n_kernel_launches = r.n_kernel_launches

    def close(self):
        if self.res is not None:
            self.ctx.result_free(self.res)
            self.res = None
        if self.tree is not None:
            self.ctx.tree_free(self.tree)
            self.tree = None

    # -- look-ups
    def dir_id(self, path) -> int:
        """Directory id of a Walk path (exact string first, then cleaned)."""
        b = os.fsencode(path)
        d = self.path_to_dir.get(b)
        if d is None:
            d = self.path_to_dir.get(lib.path_clean(b))
        return -1 if d is None else d

    def entry_path(self, e: int) -> bytes:
        d = self.entry_dir[e]
        base = lib.path_clean(self.broot) if d == 0 else self.dir_path[d]
        if base == b".":
            return self.entry_name[e]
        return base + (b"" if base.endswith(b"/") else b"/") + self.entry_name[e]

    def file_path(self, f: int) -> bytes:
        return self.entry_path(self.file_entry[f])

    def entries_with(self, feature_bit: int):
        """Entries carrying a name-feature bit, in filepath.Walk order (GetFilesByName / GetFilesByExt)."""
        return lib.list_entries(self.tree, feature_bit, self.res.contents.entry_feat)

    def dockerfiles(self):
        """Paths isDockerFile() accepts, in filepath.Walk order (dockerfile2kube.go:154-167).  File ids
        are numbered by (directory id, name), which is not Walk order when a file's name sorts after a
        sub-directory of the same directory: the entries go through the same ordering as
        GetFilesByName / GetFilesByExt."""
        feat = (lib.C.c_uint32 * max(1, self.n_entries))()
        any_set = False
        for f in range(self.n_files):
            if self.file_flags[f] & _abi.FILE_DOCKERFILE:
                feat[self.file_entry[f]] = 1
                any_set = True
        if not any_set:
            return []
        return [self.entry_path(e) for e in lib.list_entries(self.tree, 0, feat)]

    def option_json(self, d: int, option: int):
        """What the accepting built-in detect script `option` (a DirOpts bit number: 0-11 Dockerfile options,
        12-17 S2I) prints for directory d, from this pass's arrays -- nothing is executed (m2k_option_json).
        None = the option does not accept the directory."""
        return lib.option_json(self.tree, self.res, d, option)

    def main_script_rel_path(self, d: int) -> str:
        """What the python detect scripts print as main_script_rel_path / app_file (a11)."""
        f = self.dir_main_file[d]
        if f == _abi.NONE:
            return ""
        return os.fsdecode(os.path.relpath(self.file_path(f), self.dir_path[d]))


def detect(root: str, pre=(), flavor=None, skip_content=False, cache=False) -> Detection:
    """One detection pass.  The reference re-walks the tree on every call, so nothing is cached unless
    the caller asks for it (cache=True keeps the Detection, and the host memory it holds, until reset();
    a cached answer goes stale when the tree changes).  Without the cache the caller closes it."""
    key = (root, tuple(pre), sh_flavor() if flavor is None else flavor, skip_content)
    if cache and key in _cache:
        return _cache[key]
    d = Detection(root, pre, flavor, skip_content)
    if cache:
        _cache[key] = d
    return d
