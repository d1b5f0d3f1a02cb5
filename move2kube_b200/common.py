"""Mirror of the three internal/common helpers on the path (reference: internal/common/utils.go:47-82,
85-120, 308-315), answered from the packed tree instead of one more filesystem walk each."""
import os
import stat

from . import _abi, detection, lib

IgnoreFilename = ".m2kignore"

_NAME_BITS = {".m2kignore": _abi.N_IGNOREFILE, "m2kdfdetect.sh": _abi.N_USER_DF, "m2ks2idetect.sh": _abi.N_USER_S2I}


def IsStringPresent(lst, value: str) -> bool:
    """utils.go:308-315: linear scan with strings.EqualFold (Unicode simple case folding)."""
    v = os.fsencode(value)
    return any(lib.equal_fold(os.fsencode(x), v) for x in lst)


def _root_is_not_a_directory(inputPath: str) -> bool:
    """filepath.Walk Lstats the root: anything but a real directory (a file, a symlink -- even one to
    a directory) is reported once, as itself, and nothing below it is looked at (utils_test.go:41-50, 98-107)."""
    return not stat.S_ISDIR(os.lstat(inputPath).st_mode)


def GetFilesByName(inputPath: str, names):
    """utils.go:85-120.  Returns (files, err): non-directories whose basename equals one of `names`,
    in filepath.Walk order; a missing root is an error."""
    if not os.path.lexists(inputPath):
        return [], FileNotFoundError(inputPath)
    if _root_is_not_a_directory(inputPath):
        return ([inputPath] if os.path.basename(inputPath) in names else []), None
    det = detection.detect(inputPath, skip_content=True)
    try:
        return _files_by_name(det, names), None
    finally:
        det.close()


def _files_by_name(det, names):
    out = []
    known = [n for n in names if n in _NAME_BITS]
    other = [os.fsencode(n) for n in names if n not in _NAME_BITS]
    if known and not other and len(known) == 1:
        return [os.fsdecode(det.entry_path(e)) for e in det.entries_with(_NAME_BITS[known[0]])]
    # general case: any basename -- filter the packed name table (still no filesystem walk)
    t = det.tree.contents
    want = set(os.fsencode(n) for n in names)
    feat = (lib.C.c_uint32 * max(1, det.n_entries))()
    for e in range(det.n_entries):
        if det.entry_name[e] in want and (t.entry_type[e] & 0xF) not in (_abi.T_DIR, _abi.T_ERR):
            feat[e] = 1
    for e in lib.list_entries(det.tree, 0, feat):
        out.append(os.fsdecode(det.entry_path(e)))
    return out


def GetFilesByExt(inputPath: str, exts):
    """utils.go:47-82.  filepath.Ext of the basename must equal one of `exts`."""
    if not os.path.lexists(inputPath):
        return None, FileNotFoundError(inputPath)
    if _root_is_not_a_directory(inputPath):
        base = os.path.basename(inputPath)
        i = base.rfind(".")  # filepath.Ext: from the last dot of the base name, a leading dot included
        return ([inputPath] if i >= 0 and base[i:] in exts else []), None
    det = detection.detect(inputPath, skip_content=True)
    try:
        return _files_by_ext(det, exts), None
    finally:
        det.close()


def _files_by_ext(det, exts):
    if sorted(exts) == [".yaml", ".yml"]:
        return [os.fsdecode(det.entry_path(e)) for e in det.entries_with(_abi.N_YAML)]
    t = det.tree.contents
    feat = (lib.C.c_uint32 * max(1, det.n_entries))()
    bexts = [os.fsencode(x) for x in exts]
    for e in range(det.n_entries):
        if (t.entry_type[e] & 0xF) in (_abi.T_DIR, _abi.T_ERR):
            continue
        nm = det.entry_name[e]
        i = nm.rfind(b".")
        ext = nm[i:] if i >= 0 else b""
        if ext in bexts:
            feat[e] = 1
    return [os.fsdecode(det.entry_path(e)) for e in lib.list_entries(det.tree, 0, feat)]
