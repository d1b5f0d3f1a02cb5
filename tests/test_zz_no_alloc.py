"""GPU test of m2k_scan_flags_noalloc, the byte scan for hosts that cannot allocate GPU memory at the
point where they call the library (work being captured or replayed, a locked device-memory arena, a
fixed quota).  The caller -- here the test, through torch -- owns every buffer; the library brings
only its built-in record table.

The file is named so that pytest collects it last: if it fails, a run that stops at the first
failure has already run everything else.  It is checked against the existing entry point
(m2k_detect) AND against the CPU oracle's per-file flags on two calls in a row, the second on a
smaller and different tree, so records left in the shared table by the first call cannot leak into
the second answer.  Bit-exact, like every other test here.  "Allocates nothing" is checked with the
driver's own count of free device memory (cudaMemGetInfo), which sees a cudaMalloc made by any
library -- torch's allocator statistics would not.
"""
import ctypes as C

import pytest

from move2kube_b200 import _abi, lib

pytestmark = pytest.mark.gpu

CHUNK = 65536


def _params(seed, n_files, n_dirs, total):
    return _abi.SynthParams(seed=seed, n_files=n_files, n_dirs=n_dirs, max_depth=5, n_ignore_files=3, total_bytes=total)


def _run_noalloc(torch, L, tree):
    """Copies the tree's content into caller-owned device buffers and calls the no-allocation entry point."""
    t = tree.contents
    nf = t.n_files
    lens = t.file_len[0:nf]
    offs = t.file_off[0:nf]
    first = [0]
    for ln in lens:
        first.append(first[-1] + (ln + CHUNK - 1) // CHUNK)
    host = torch.frombuffer(bytearray(C.string_at(t.bytes, t.n_bytes)), dtype=torch.uint8)
    d_bytes = host.cuda()
    d_off = torch.tensor(offs, dtype=torch.int64).cuda()
    d_len = torch.tensor(lens, dtype=torch.int64).cuda()
    d_first = torch.tensor(first, dtype=torch.int32).cuda()
    d_flags = torch.full((max(nf, 1),), 0xEE, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    # the caller's stream, not the default one; the first call on the device (the set-up call) is made by
    # whichever test reaches this first, outside any measurement
    stream = torch.cuda.Stream()
    rc = L.m2k_scan_flags_noalloc(0, stream.cuda_stream, d_bytes.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), d_first.data_ptr(),
                                  nf, first[-1], d_flags.data_ptr())
    assert rc == 0, L.m2k_global_error()
    stream.synchronize()
    d_flags.fill_(0xEE)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    dev_before = torch.cuda.current_device()
    rc = L.m2k_scan_flags_noalloc(0, stream.cuda_stream, d_bytes.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), d_first.data_ptr(),
                                  nf, first[-1], d_flags.data_ptr())
    assert rc == 0, L.m2k_global_error()
    stream.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the device's free memory"
    assert torch.cuda.current_device() == dev_before
    return d_flags.cpu().tolist()[:nf], first[-1]


def _oracle_flags(tree):
    """HAS_MAIN | LONG_LINE of every file from the CPU oracle (memmem / line lengths over the host bytes)."""
    import oracle_lib
    t = tree.contents
    out = []
    for f in range(t.n_files):
        data = C.string_at(C.addressof(t.bytes.contents) + t.file_off[f], t.file_len[f]) if t.file_len[f] else b""
        out.append(oracle_lib.file_flags(data) & (_abi.FILE_HAS_MAIN | _abi.FILE_LONG_LINE))
    return out


def test_no_alloc_scan_matches_the_allocating_path_twice_in_a_row():
    import torch
    L = lib.load()
    ctx = lib.Context(0)
    try:
        # first a tree with multi-chunk files, lines over 64 KiB and "__main__" hits; then a smaller, different one
        for seed, n_files, n_dirs, total in ((21, 1800, 100, 80 << 20), (22, 700, 60, 12 << 20)):
            tree = ctx.synth_tree(_params(seed, n_files, n_dirs, total), host_bytes=True)
            res = ctx.detect(tree)
            want = [x & (_abi.FILE_HAS_MAIN | _abi.FILE_LONG_LINE) for x in res.contents.file_flags[0:res.contents.n_files]]
            got, n_chunks = _run_noalloc(torch, L, tree)
            assert got == want
            assert got == _oracle_flags(tree)
            assert any(x & _abi.FILE_HAS_MAIN for x in want)
            if seed == 21:
                assert n_chunks > n_files and any(x & _abi.FILE_LONG_LINE for x in want)
            ctx.result_free(res)
            ctx.tree_free(tree)
    finally:
        ctx.close()


def test_no_alloc_scan_refuses_what_does_not_fit():
    import torch
    L = lib.load()
    z = torch.zeros(64, dtype=torch.uint8, device="cuda")
    rc = L.m2k_scan_flags_noalloc(0, None, z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), 1, 262145, z.data_ptr())
    assert rc == _abi.ERR_LIMIT
    msg = L.m2k_global_error().decode()
    assert "262144" in msg and "16 GiB" in msg
    assert L.m2k_scan_flags_noalloc(0, None, None, z.data_ptr(), z.data_ptr(), z.data_ptr(), 1, 1, z.data_ptr()) == _abi.ERR_ARG
