import sys, os, time, ctypes as C
sys.path.insert(0, "/root/repo")
import bench
from move2kube_b200 import lib, _abi
ctx = lib.Context(0)
for w in ("B", "P"):
    p = bench.synth_params(w)
    tree = ctx.synth_tree(p, host_bytes=True)   # pinned host bytes
    for i in range(3):
        t0 = time.perf_counter(); res = ctx.detect(tree); t1 = time.perf_counter()
        ctx.result_free(res); t2 = time.perf_counter()
        print(w, "detect %.2f ms  result_free %.2f ms  ms_h2d %.2f" % (1e3*(t1-t0), 1e3*(t2-t1), 0), flush=True)
    ctx.tree_free(tree)
