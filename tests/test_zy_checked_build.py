"""Runs near the end of the GPU tier (before the no-allocation test): when the library under test is the
CHECKED build (tools/scan_variants.py "checked": -DM2K_CHECKED=1, selected with M2K_LIB=...), every kernel access
that would have fallen outside its buffer during ALL the tests before this one was counted instead of made.  The
counters must be zero.  This is the repository's own stand-in for a memory checker (compute-sanitizer is not
available on the GPU pool: profiles/r2_compute_sanitizer_closed.log).  With the product build the test only checks
that the diagnostic entry point answers."""
import pytest

from move2kube_b200 import _abi, lib

pytestmark = pytest.mark.gpu


def test_no_out_of_bounds_access_was_counted():
    ctx = lib.Context(0)
    try:
        # make sure the counters saw at least one full pass in this process even when run alone
        p = _abi.SynthParams(seed=5, n_files=900, n_dirs=60, max_depth=5, n_ignore_files=3, total_bytes=24 << 20,
                             n_huge=1, huge_bytes=3 << 20)
        tree = ctx.synth_tree(p, host_bytes=True)
        res = ctx.detect(tree)
        ctx.result_free(res)
        ctx.tree_free(tree)
        checked, counters = ctx.debug_violations()
        assert counters == [0] * 8, "out-of-bounds accesses counted by the checked build: %r" % (counters,)
        print("checked build: %s" % checked)
    finally:
        ctx.close()
