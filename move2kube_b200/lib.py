"""ctypes binding of libm2kdetect.so (include/m2kdetect.h).

The shared library is the product; this module only loads it and gives the calls Python
signatures.  It fails loudly when the library is missing or a call fails: there is no CPU path.
"""
import ctypes as C
import os

from . import _abi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("M2K_LIB") or os.path.join(HERE, "libm2kdetect.so")  # M2K_LIB: a tuning variant

_lib = None


class M2KError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("libm2kdetect error %d: %s" % (code, message))
        self.code = code
        self.message = message


# every symbol include/m2kdetect.h declares
EXPORTS = [
    "m2k_abi_version", "m2k_global_error", "m2k_ctx_create", "m2k_ctx_destroy", "m2k_last_error", "m2k_device_info",
    "m2k_pack_tree", "m2k_tree_free", "m2k_pinned_alloc", "m2k_pinned_free", "m2k_synth_tree", "m2k_synth_fill_host", "m2k_detect", "m2k_result_free",
    "m2k_upload", "m2k_synth_fill_device", "m2k_scan", "m2k_partials", "m2k_merge_partials", "m2k_file_flags_get", "m2k_file_flags_set", "m2k_prepare", "m2k_finalize",
    "m2k_devtree_free", "m2k_scan_kernel_ms", "m2k_phase_ms", "m2k_launch_count", "m2k_stream", "m2k_path_clean", "m2k_equal_fold",
    "m2k_list_entries", "m2k_scan_flags_noalloc", "m2k_host_prepare_probe",
    "m2k_chunk_count", "m2k_plan_shards", "m2k_upload_chunks", "m2k_pieces", "m2k_merge_pieces",
    "m2k_chunk_locate", "m2k_file_flag", "m2k_settle_file", "m2k_devtree_content_bytes", "m2k_root_extra_bytes",
    "m2k_export_async", "m2k_import_async", "m2k_pack_detect", "m2k_is_dockerfile",
    "m2k_exchange_open", "m2k_exchange_attach", "m2k_exchange_attach_local", "m2k_exchange_collect", "m2k_exchange_wait", "m2k_exchange_wait_previous", "m2k_reupload", "m2k_detect_multi", "m2k_option_json", "m2k_ctx_trim", "m2k_debug_violations",
]


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: build it with `python -m move2kube_b200.build` (nvcc, sm_100a). "
                          "move2kube_b200 has no fallback implementation." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, TP, RP = C.c_void_p, C.POINTER(_abi.Tree), C.POINTER(_abi.Result)
    PP = C.POINTER(_abi.DetectParams)
    L.m2k_abi_version.restype = C.c_int
    L.m2k_global_error.restype = C.c_char_p
    L.m2k_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.m2k_ctx_destroy.argtypes = [vp]
    L.m2k_ctx_destroy.restype = None
    L.m2k_last_error.argtypes = [vp]
    L.m2k_last_error.restype = C.c_char_p
    L.m2k_device_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_uint64), C.c_char_p, C.c_size_t]
    L.m2k_pack_tree.argtypes = [vp, C.c_char_p, C.c_int, C.POINTER(TP)]
    L.m2k_tree_free.argtypes = [vp, TP]
    L.m2k_tree_free.restype = None
    L.m2k_pinned_alloc.argtypes = [vp, C.c_size_t]
    L.m2k_pinned_alloc.restype = vp
    L.m2k_pinned_free.argtypes = [vp, vp]
    L.m2k_pinned_free.restype = None
    L.m2k_synth_tree.argtypes = [vp, C.POINTER(_abi.SynthParams), C.c_int, C.POINTER(TP)]
    L.m2k_synth_fill_host.argtypes = [TP, C.c_uint64, C.c_uint32, C.c_uint32, vp, C.c_int]
    L.m2k_detect.argtypes = [vp, TP, PP, C.POINTER(RP)]
    L.m2k_result_free.argtypes = [RP]
    L.m2k_result_free.restype = None
    L.m2k_upload.argtypes = [vp, TP, C.c_uint32, C.c_uint32, C.POINTER(vp)]
    L.m2k_synth_fill_device.argtypes = [vp, vp, C.c_uint64]
    L.m2k_scan.argtypes = [vp, vp, PP]
    L.m2k_partials.argtypes = [vp, vp, vp, vp]
    L.m2k_merge_partials.argtypes = [vp, vp, vp, vp, C.c_uint32]
    L.m2k_file_flags_get.argtypes = [vp, vp, vp]
    L.m2k_file_flags_set.argtypes = [vp, vp, vp]
    L.m2k_prepare.argtypes = [vp, vp, PP]
    L.m2k_finalize.argtypes = [vp, vp, PP, C.POINTER(RP)]
    L.m2k_devtree_free.argtypes = [vp, vp]
    L.m2k_devtree_free.restype = None
    L.m2k_scan_kernel_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.m2k_phase_ms.argtypes = [vp, C.POINTER(C.c_float), C.c_int]
    L.m2k_launch_count.argtypes = [vp]
    L.m2k_launch_count.restype = C.c_uint64
    L.m2k_stream.argtypes = [vp]
    L.m2k_stream.restype = vp
    L.m2k_path_clean.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
    L.m2k_equal_fold.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    L.m2k_list_entries.argtypes = [TP, C.c_uint32, _abi.u32p, _abi.u32p, C.c_uint32]
    L.m2k_scan_flags_noalloc.argtypes = [C.c_int, vp, vp, vp, vp, vp, C.c_uint32, C.c_uint32, vp]
    L.m2k_chunk_count.argtypes = [TP]
    L.m2k_chunk_count.restype = C.c_uint64
    L.m2k_plan_shards.argtypes = [TP, C.c_uint32, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
    L.m2k_root_extra_bytes.argtypes = [TP]
    L.m2k_root_extra_bytes.restype = C.c_uint64
    L.m2k_upload_chunks.argtypes = [vp, TP, C.c_uint64, C.c_uint64, C.POINTER(vp)]
    L.m2k_pieces.argtypes = [vp, vp, vp]
    L.m2k_merge_pieces.argtypes = [vp, vp, vp, C.c_uint32]
    L.m2k_chunk_locate.argtypes = [TP, C.c_uint64, _abi.u32p, _abi.u32p]
    L.m2k_file_flag.argtypes = [vp, vp, C.c_uint32, C.POINTER(C.c_uint8)]
    L.m2k_settle_file.argtypes = [vp, vp, C.c_uint32, C.c_uint32]
    L.m2k_devtree_content_bytes.argtypes = [vp]
    L.m2k_devtree_content_bytes.restype = C.c_uint64
    L.m2k_export_async.argtypes = [vp, vp, vp, vp, vp, vp]
    L.m2k_import_async.argtypes = [vp, vp, vp, vp, vp, vp, C.c_uint32]
    L.m2k_pack_detect.argtypes = [vp, C.c_char_p, PP, C.POINTER(TP), C.POINTER(RP)]
    L.m2k_reupload.argtypes = [vp, vp, TP]
    L.m2k_ctx_trim.argtypes = [vp]
    L.m2k_debug_violations.argtypes = [vp, C.POINTER(C.c_uint64)]
    L.m2k_option_json.argtypes = [TP, RP, C.c_uint32, C.c_uint32, C.c_char_p, C.c_size_t]
    L.m2k_detect_multi.argtypes = [C.POINTER(vp), C.c_uint32, TP, PP, C.POINTER(RP)]
    L.m2k_exchange_open.argtypes = [vp, vp, C.c_uint32, vp]
    L.m2k_exchange_attach.argtypes = [vp, vp, C.c_uint32, vp]
    L.m2k_exchange_attach_local.argtypes = [vp, vp, C.c_uint32, vp, vp]
    L.m2k_exchange_collect.argtypes = [vp, vp]
    L.m2k_exchange_wait.argtypes = [vp, vp]
    L.m2k_exchange_wait_previous.argtypes = [vp, vp]
    L.m2k_is_dockerfile.argtypes = [vp, C.c_char_p, C.c_uint64, _abi.u32p]
    L.m2k_host_prepare_probe.argtypes = [TP, PP, C.c_uint32, C.POINTER(C.c_double), _abi.u32p, _abi.u32p]
    for name in EXPORTS:
        fn = getattr(L, name)
        if fn.restype is C.c_int and name not in ("m2k_abi_version", "m2k_equal_fold", "m2k_list_entries"):
            pass
    _lib = L
    return L


def _check(rc, ctx=None):
    if rc != _abi.OK:
        L = load()
        msg = L.m2k_last_error(ctx) if ctx else L.m2k_global_error()
        raise M2KError(rc, (msg or b"").decode("utf-8", "replace"))


def make_params(flavor=_abi.SH_BASH, pre=(), skip_content=False):
    """Returns (DetectParams, keepalive)."""
    arr = (C.c_char_p * max(1, len(pre)))(*[bytes(p) for p in pre])
    p = _abi.DetectParams(flavor, len(pre), arr, 1 if skip_content else 0, 0)
    return p, arr


class Context:
    """One CUDA device (m2k_ctx).  Creating it without a GPU raises M2KError(M2K_ERR_CUDA)."""

    def __init__(self, device=0):
        self.L = load()
        h = C.c_void_p()
        _check(self.L.m2k_ctx_create(device, C.byref(h)))
        self.h = h
        self.device = device

    def close(self):
        if self.h:
            self.L.m2k_ctx_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def debug_violations(self):
        """(checked build?, [8 counters]) -- see m2k_debug_violations."""
        out = (C.c_uint64 * 8)()
        rc = self.L.m2k_debug_violations(self.h, out)
        if rc < 0:
            _check(rc, self.h)
        return bool(rc), list(out)

    def trim(self):
        """Give the device buffers the context keeps between calls back to the driver."""
        _check(self.L.m2k_ctx_trim(self.h), self.h)

    def device_info(self):
        sm, mem, name = C.c_int(), C.c_uint64(), C.create_string_buffer(256)
        _check(self.L.m2k_device_info(self.h, C.byref(sm), C.byref(mem), name, 256), self.h)
        return {"sm_count": sm.value, "hbm_bytes": mem.value, "name": name.value.decode()}

    # -- host trees
    def pack_tree(self, root: bytes, read_content=True):
        tp = C.POINTER(_abi.Tree)()
        _check(self.L.m2k_pack_tree(self.h, root, 1 if read_content else 0, C.byref(tp)), self.h)
        return tp

    def synth_tree(self, params: _abi.SynthParams, host_bytes: bool):
        tp = C.POINTER(_abi.Tree)()
        _check(self.L.m2k_synth_tree(self.h, C.byref(params), 1 if host_bytes else 0, C.byref(tp)), self.h)
        return tp

    def pinned_alloc(self, nbytes):
        p = self.L.m2k_pinned_alloc(self.h, nbytes)
        if not p:
            _check(_abi.ERR_NOMEM, self.h)
        return p

    def pinned_free(self, p):
        self.L.m2k_pinned_free(self.h, p)

    def tree_free(self, tree):
        self.L.m2k_tree_free(self.h, tree)

    # -- whole path
    def detect(self, tree, flavor=_abi.SH_BASH, pre=(), skip_content=False):
        p, keep = make_params(flavor, pre, skip_content)
        rp = C.POINTER(_abi.Result)()
        _check(self.L.m2k_detect(self.h, tree, C.byref(p), C.byref(rp)), self.h)
        return rp

    def pack_detect(self, root: bytes, flavor=_abi.SH_BASH, pre=(), skip_content=False):
        """(tree without bytes, result): pack and detect in one call, bytes streamed from the files."""
        p, keep = make_params(flavor, pre, skip_content)
        tp = C.POINTER(_abi.Tree)()
        rp = C.POINTER(_abi.Result)()
        _check(self.L.m2k_pack_detect(self.h, root, C.byref(p), C.byref(tp), C.byref(rp)), self.h)
        return tp, rp

    def result_free(self, res):
        self.L.m2k_result_free(res)

    def is_dockerfile(self, data: bytes) -> bool:
        """isDockerFile (dockerfile2kube.go:117-144) for one file's bytes, through the same kernels."""
        flags = C.c_uint32()
        _check(self.L.m2k_is_dockerfile(self.h, data, len(data), C.byref(flags)), self.h)
        return bool(flags.value & _abi.FILE_DOCKERFILE)

    # -- staged
    def upload(self, tree, file_begin=0, file_end=None):
        if file_end is None:
            file_end = tree.contents.n_files
        h = C.c_void_p()
        _check(self.L.m2k_upload(self.h, tree, file_begin, file_end, C.byref(h)), self.h)
        return h

    def reupload(self, dt, tree):
        _check(self.L.m2k_reupload(self.h, dt, tree), self.h)

    def upload_chunks(self, tree, chunk_begin, chunk_end):
        h = C.c_void_p()
        _check(self.L.m2k_upload_chunks(self.h, tree, chunk_begin, chunk_end, C.byref(h)), self.h)
        return h

    def pieces(self, dt, dst_ptr):
        _check(self.L.m2k_pieces(self.h, dt, dst_ptr), self.h)

    def merge_pieces(self, dt, all_pieces_ptr, n_ranks):
        _check(self.L.m2k_merge_pieces(self.h, dt, all_pieces_ptr, n_ranks), self.h)

    def export_async(self, dt, content_ptr, main_ptr, pieces_ptr, flags_ptr):
        _check(self.L.m2k_export_async(self.h, dt, content_ptr, main_ptr, pieces_ptr, flags_ptr), self.h)

    def import_async(self, dt, all_content_ptr, all_main_ptr, all_pieces_ptr, flags_ptr, n_ranks):
        _check(self.L.m2k_import_async(self.h, dt, all_content_ptr, all_main_ptr, all_pieces_ptr, flags_ptr, n_ranks), self.h)

    # -- exchange over peer memory
    def exchange_open(self, dt, n_ranks) -> bytes:
        """Root: returns the 192-byte handle the other ranks attach with."""
        buf = C.create_string_buffer(192)
        _check(self.L.m2k_exchange_open(self.h, dt, n_ranks, buf), self.h)
        return buf.raw

    def exchange_attach(self, dt, rank, handle: bytes):
        buf = C.create_string_buffer(bytes(handle), 192)
        _check(self.L.m2k_exchange_attach(self.h, dt, rank, buf), self.h)

    def exchange_attach_local(self, dt, rank, root_ctx, root_dt):
        _check(self.L.m2k_exchange_attach_local(self.h, dt, rank, root_ctx.h, root_dt), self.h)

    def exchange_collect(self, dt):
        _check(self.L.m2k_exchange_collect(self.h, dt), self.h)

    def exchange_wait(self, dt):
        _check(self.L.m2k_exchange_wait(self.h, dt), self.h)

    def exchange_wait_previous(self, dt):
        _check(self.L.m2k_exchange_wait_previous(self.h, dt), self.h)

    def shard_content_bytes(self, dt) -> int:
        return self.L.m2k_devtree_content_bytes(dt)

    def file_flag(self, dt, f) -> int:
        out = C.c_uint8()
        _check(self.L.m2k_file_flag(self.h, dt, f, C.byref(out)), self.h)
        return out.value

    def settle_file(self, dt, f, flags):
        _check(self.L.m2k_settle_file(self.h, dt, f, flags), self.h)

    def synth_fill_device(self, dt, seed):
        _check(self.L.m2k_synth_fill_device(self.h, dt, seed), self.h)

    def scan(self, dt, flavor=_abi.SH_BASH, skip_content=False):
        p, keep = make_params(flavor, (), skip_content)
        _check(self.L.m2k_scan(self.h, dt, C.byref(p)), self.h)

    def partials(self, dt, dst_content_ptr, dst_main_ptr):
        """Copies this shard's per-directory partials into caller-owned device buffers."""
        _check(self.L.m2k_partials(self.h, dt, dst_content_ptr, dst_main_ptr), self.h)

    def merge_partials(self, dt, all_content_ptr, all_main_ptr, n_ranks):
        _check(self.L.m2k_merge_partials(self.h, dt, all_content_ptr, all_main_ptr, n_ranks), self.h)

    def file_flags_get(self, dt, dst_ptr):
        _check(self.L.m2k_file_flags_get(self.h, dt, dst_ptr), self.h)

    def file_flags_set(self, dt, src_ptr):
        _check(self.L.m2k_file_flags_set(self.h, dt, src_ptr), self.h)

    def prepare(self, dt, flavor=_abi.SH_BASH, pre=()):
        p, keep = make_params(flavor, pre, False)
        _check(self.L.m2k_prepare(self.h, dt, C.byref(p)), self.h)

    def finalize(self, dt, flavor=_abi.SH_BASH, pre=()):
        p, keep = make_params(flavor, pre, False)
        rp = C.POINTER(_abi.Result)()
        _check(self.L.m2k_finalize(self.h, dt, C.byref(p), C.byref(rp)), self.h)
        return rp

    def devtree_free(self, dt):
        self.L.m2k_devtree_free(self.h, dt)

    def scan_kernel_ms(self):
        ms = C.c_float()
        _check(self.L.m2k_scan_kernel_ms(self.h, C.byref(ms)), self.h)
        return ms.value

    PHASES = ["names", "scan", "combine", "dockerfile", "finalize_kernels", "host_ignore", "host_verify", "d2h"]

    def phase_ms(self):
        out = (C.c_float * 8)()
        _check(self.L.m2k_phase_ms(self.h, out, 8), self.h)
        return dict(zip(self.PHASES, [float(x) for x in out]))

    def launch_count(self):
        return int(self.L.m2k_launch_count(self.h))

    def stream(self):
        return self.L.m2k_stream(self.h)


def detect_multi(ctxs, tree, flavor=_abi.SH_BASH, pre=(), skip_content=False):
    """m2k_detect_multi: the whole path on several devices from this one process; ctxs[0] assembles the result."""
    L = load()
    p, keep = make_params(flavor, pre, skip_content)
    arr = (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])
    rp = C.POINTER(_abi.Result)()
    _check(L.m2k_detect_multi(arr, len(ctxs), tree, C.byref(p), C.byref(rp)), ctxs[0].h)
    return rp


# -- calls that need no GPU (host logic of the boundary)
def pack_tree_host(root: bytes, read_content=True):
    L = load()
    tp = C.POINTER(_abi.Tree)()
    _check(L.m2k_pack_tree(None, root, 1 if read_content else 0, C.byref(tp)))
    return tp


def synth_tree_host(params: _abi.SynthParams, host_bytes=True):
    L = load()
    tp = C.POINTER(_abi.Tree)()
    _check(L.m2k_synth_tree(None, C.byref(params), 1 if host_bytes else 0, C.byref(tp)))
    return tp


def tree_free_host(tree):
    load().m2k_tree_free(None, tree)


def chunk_count(tree) -> int:
    return load().m2k_chunk_count(tree)


def plan_shards(tree, n_ranks, split_min_bytes=64 << 20, root_extra_bytes=0):
    """n_ranks + 1 chunk cut points of near-equal cost; a file is cut only if it has at least
    `split_min_bytes` bytes; shard 0 is lighter by `root_extra_bytes` (None: the library's estimate of
    what the rank that assembles the result does besides scanning)."""
    L = load()
    if root_extra_bytes is None:
        root_extra_bytes = L.m2k_root_extra_bytes(tree)
    cuts = (C.c_uint64 * (n_ranks + 1))()
    _check(L.m2k_plan_shards(tree, n_ranks, split_min_bytes, root_extra_bytes, cuts))
    return list(cuts)


def chunk_locate(tree, chunk):
    """(file id, chunk index inside that file) of a global chunk number."""
    f, k = C.c_uint32(), C.c_uint32()
    _check(load().m2k_chunk_locate(tree, chunk, C.byref(f), C.byref(k)))
    return f.value, k.value


def option_json(tree, res, d: int, option: int):
    """Stdout of the accepting built-in detect script `option` for directory d (bytes), or None when the
    option does not accept the directory."""
    L = load()
    n = L.m2k_option_json(tree, res, d, option, None, 0)
    if n == _abi.ERR_ARG:
        return None
    if n < 0:
        raise M2KError(n, (L.m2k_global_error() or b"").decode("utf-8", "replace"))
    buf = C.create_string_buffer(n + 1)
    n = L.m2k_option_json(tree, res, d, option, buf, n + 1)
    if n < 0:
        raise M2KError(n, (L.m2k_global_error() or b"").decode("utf-8", "replace"))
    return buf.raw[:n]


def path_clean(p: bytes) -> bytes:
    buf = C.create_string_buffer(len(p) + 8)
    _check(load().m2k_path_clean(p, buf, len(p) + 8))
    return buf.value


def equal_fold(a: bytes, b: bytes) -> bool:
    return bool(load().m2k_equal_fold(a, len(a), b, len(b)))


def list_entries(tree, feature_bit, entry_feat):
    L = load()
    n = tree.contents.n_entries
    out = (C.c_uint32 * max(1, n))()
    k = L.m2k_list_entries(tree, feature_bit, entry_feat, out, n)
    if k < 0:
        raise M2KError(k, "m2k_list_entries")
    return list(out[0:k])
