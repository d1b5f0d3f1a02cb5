"""Times the host half of m2k_prepare alone (no GPU needed) on the metadata of N copies' worth of a
bench workload -- the stretch of a multi-GPU pass that only rank 0 runs.

    python tools/prepare_probe.py --workload C --scale 8 --reps 10
"""
import argparse
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import bench  # noqa: E402
from move2kube_b200 import build, lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C")
    ap.add_argument("--scale", type=int, default=1)
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    build.build()
    L = lib.load()
    p = bench.synth_params(a.workload, a.scale)
    t0 = time.time()
    tree = lib.synth_tree_host(p, host_bytes=False)
    t1 = time.time()
    ms = (C.c_double * 6)()
    nf, nr = C.c_uint32(), C.c_uint32()
    rc = L.m2k_host_prepare_probe(tree, None, a.reps, ms, C.byref(nf), C.byref(nr))
    if rc:
        raise SystemExit("probe failed: %s" % L.m2k_global_error().decode())
    t = tree.contents
    print("tree: %d files, %d dirs, %d entries (generated in %.1f s); %d rule files, %d rules" %
          (t.n_files, t.n_dirs, t.n_entries, t1 - t0, nf.value, nr.value))
    print("best of %d: all %.2f ms = order %.2f + parse %.2f + fold %.2f + table %.2f + lists %.2f" %
          ((a.reps,) + tuple(ms)))
    lib.tree_free_host(tree)


if __name__ == "__main__":
    main()
