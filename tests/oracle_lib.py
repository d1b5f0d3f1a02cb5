"""ctypes loader for the CPU oracle (oracle/libm2koracle.so).  Test infrastructure only."""
import ctypes as C
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)
from move2kube_b200 import _abi  # noqa: E402  (struct layouts only)

_lib = None


class OResult(C.Structure):
    _fields_ = [
        ("n_dirs", C.c_uint32), ("n_files", C.c_uint32), ("n_services", C.c_uint32), ("_pad", C.c_uint32),
        ("dir_opts", _abi.u32p), ("dir_state", _abi.u8p), ("dir_ignore", _abi.u8p), ("dir_main_file", _abi.u32p),
        ("service_dirs", _abi.u32p), ("file_flags", _abi.u8p),
        ("n_ignore_dirs", C.c_uint32), ("n_ignore_contents", C.c_uint32),
        ("ignore_dirs", C.POINTER(C.c_char_p)), ("ignore_contents", C.POINTER(C.c_char_p)),
    ]


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(REPO, "oracle"), "libm2koracle.so"])


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.join(REPO, "oracle", "libm2koracle.so")
    src = os.path.join(REPO, "oracle", "m2k_oracle.c")
    if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
        build()
    L = C.CDLL(path)
    TP = C.POINTER(_abi.Tree)
    L.m2ko_pack_fs.argtypes = [C.c_char_p, C.c_int, C.POINTER(TP)]
    L.m2ko_pack_fs.restype = C.c_int
    L.m2ko_tree_free.argtypes = [TP]
    L.m2ko_tree_free.restype = None
    L.m2ko_detect.argtypes = [TP, C.c_int, C.c_uint32, C.POINTER(C.c_char_p), C.c_int, C.c_int,
                              C.POINTER(C.POINTER(OResult))]
    L.m2ko_detect.restype = C.c_int
    L.m2ko_result_free.argtypes = [C.POINTER(OResult)]
    L.m2ko_result_free.restype = None
    L.m2ko_dir_path.argtypes = [TP, C.c_uint32]
    L.m2ko_dir_path.restype = C.c_void_p
    L.m2ko_free.argtypes = [C.c_void_p]
    L.m2ko_free.restype = None
    L.m2ko_file_flags.argtypes = [C.c_char_p, C.c_uint64]
    L.m2ko_file_flags.restype = C.c_uint32
    L.m2ko_dockerfile_flags.argtypes = [C.c_char_p, C.c_uint64]
    L.m2ko_dockerfile_flags.restype = C.c_uint32
    L.m2ko_scan_files.argtypes = [TP, C.c_uint32, C.c_uint32, _abi.u8p, C.c_int]
    L.m2ko_scan_files.restype = C.c_int
    L.m2ko_equal_fold.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    L.m2ko_equal_fold.restype = C.c_int
    L.m2ko_clean.argtypes = [C.c_char_p, C.c_size_t]
    L.m2ko_clean.restype = C.c_void_p
    L.m2ko_join.argtypes = [C.c_char_p, C.c_char_p]
    L.m2ko_join.restype = C.c_void_p
    L.m2ko_files_by_name.argtypes = [TP, C.c_char_p, _abi.u32p, C.c_uint32]
    L.m2ko_files_by_name.restype = C.c_uint32
    L.m2ko_dir_options.argtypes = [TP, C.c_uint32, C.c_int]
    L.m2ko_dir_options.restype = C.c_uint32
    _lib = L
    return L


def _take_str(ptr):
    s = C.string_at(ptr)
    lib().m2ko_free(ptr)
    return s


def clean(p: bytes) -> bytes:
    return _take_str(lib().m2ko_clean(p, len(p)))


def join(a: bytes, b: bytes) -> bytes:
    return _take_str(lib().m2ko_join(a, b))


def equal_fold(a: bytes, b: bytes) -> bool:
    return bool(lib().m2ko_equal_fold(a, len(a), b, len(b)))


def dir_path(tree, d: int) -> bytes:
    return _take_str(lib().m2ko_dir_path(tree, d))


def pack_fs(root: bytes, read_content=True):
    tp = C.POINTER(_abi.Tree)()
    rc = lib().m2ko_pack_fs(root, 1 if read_content else 0, C.byref(tp))
    return rc, tp


def detect(tree, flavor=_abi.SH_BASH, pre=(), scan_content=True, n_threads=1):
    arr = (C.c_char_p * max(1, len(pre)))(*pre)
    rp = C.POINTER(OResult)()
    rc = lib().m2ko_detect(tree, flavor, len(pre), arr, 1 if scan_content else 0, n_threads, C.byref(rp))
    assert rc == 0
    return rp


def file_flags(data: bytes) -> int:
    return lib().m2ko_file_flags(data, len(data))


def dockerfile_flags(data: bytes) -> int:
    return lib().m2ko_dockerfile_flags(data, len(data))
