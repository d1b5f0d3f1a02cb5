#!/usr/bin/env python3
"""Builds tuning variants of the byte-scan kernel (different -D values of kernels.cu's tunables) into
build/variants/ and, with --run on a GPU box, times each with bench.py (config C, kernel-only).

    python tools/scan_variants.py --build          # here (nvcc cross-compiles)
    python tools/scan_variants.py --run            # on the B200 (via gpurun)
"""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
OUTDIR = os.path.join(REPO, "move2kube_b200", "variants")

VARIANTS = {
    "base": [],
    "w32": ["M2K_SCAN_WAVES=32"],
    "w128": ["M2K_SCAN_WAVES=128"],
    "w512": ["M2K_SCAN_WAVES=512"],
    "w4096": ["M2K_SCAN_WAVES=4096"],
    "w4096_t128": ["M2K_SCAN_WAVES=4096", "M2K_SCAN_THREADS=128", "M2K_SCAN_MINBLOCKS=8"],
    "w4096_t64": ["M2K_SCAN_WAVES=8192", "M2K_SCAN_THREADS=64", "M2K_SCAN_MINBLOCKS=16"],
    "w128_rows4_b6": ["M2K_SCAN_WAVES=128", "M2K_SCAN_ROWS=4", "M2K_SCAN_MINBLOCKS=6"],
}


def main():
    from move2kube_b200 import build as b
    os.makedirs(OUTDIR, exist_ok=True)
    if "--build" in sys.argv:
        for name, defs in VARIANTS.items():
            b.build(force=True, defines=defs, out=os.path.join(OUTDIR, "libm2kdetect_%s.so" % name))
            print("built", name)
    if "--run" in sys.argv:
        wl = "C"
        for name in VARIANTS:
            env = dict(os.environ, M2K_LIB=os.path.join(OUTDIR, "libm2kdetect_%s.so" % name))
            p = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--workload", wl, "--steps", "5", "--warmup", "3",
                                "--no-e2e", "--no-cpu"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
            try:
                j = json.loads(p.stdout.strip().splitlines()[-1])
                print("%-12s scan %.3f ms  %.0f GB/s  frac %.3f  step %.2f ms" % (name, j["roofline"]["kernel_ms"], j["roofline"]["achieved"],
                                                                                j["roofline"]["frac"], j["ms_per_step"]), flush=True)
            except Exception as e:
                print(name, "FAILED", e, p.stderr[-300:], flush=True)


if __name__ == "__main__":
    main()
