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
    # round 2: work items drawn from a counter instead of strided; the bulk-copy engine (cp.async.bulk + mbarrier
    # rings in shared memory) feeding the same scan
    "quota16": ["M2K_SCAN_QUOTA_WAVES=16"],
    "quota32": ["M2K_SCAN_QUOTA_WAVES=32"],
    "quota64": ["M2K_SCAN_QUOTA_WAVES=64"],
    "mb3": ["M2K_SCAN_MINBLOCKS=3"],
    "prefetch": ["M2K_SCAN_PREFETCH=1"],  # L2 prefetch of the next batch while the current one is processed: C 14.10 vs 13.75 ms, B 0.159 vs 0.163, 32-512 MB trees unchanged
    "checked": ["M2K_CHECKED=1"],  # own bounds checks (m2k_debug_violations), run under the parity tests
    "strided": ["M2K_SCAN_TICKET=0"],  # round 1's distribution: item += number of warps, 32 waves of blocks
    "bulk_w16_s3": ["M2K_SCAN_BULK=1", "M2K_BULK_WARPS=16", "M2K_BULK_STAGES=3"],
    "bulk_w24_s2": ["M2K_SCAN_BULK=1", "M2K_BULK_WARPS=24", "M2K_BULK_STAGES=2"],
    "bulk_w12_s4": ["M2K_SCAN_BULK=1", "M2K_BULK_WARPS=12", "M2K_BULK_STAGES=4"],
    "bulk_w16_s6_r4": ["M2K_SCAN_BULK=1", "M2K_BULK_WARPS=16", "M2K_BULK_STAGES=6", "M2K_SCAN_ROWS=4"],
    "bulk_w32_s3_r4": ["M2K_SCAN_BULK=1", "M2K_BULK_WARPS=32", "M2K_BULK_STAGES=3", "M2K_SCAN_ROWS=4"],
    # Dockerfile body stripes (the product uses 128 bytes)
    "df_stripe256": ["M2K_DF_STRIPE=256"],  # B: dockerfile phase 0.165 ms (256) / 0.116 (128) / 0.087 (64); C: 0.392 / 0.383 / 0.419
    "df_stripe64": ["M2K_DF_STRIPE=64"],
    # round 1 (kept for the record; "base" won): M2K_SCAN_WAVES 32/128/512/4096, 128- and 64-thread blocks, 4 rows
}


def main():
    from move2kube_b200 import build as b
    os.makedirs(OUTDIR, exist_ok=True)
    if "--build" in sys.argv:
        only = sys.argv[sys.argv.index("--only") + 1].split(",") if "--only" in sys.argv else list(VARIANTS)
        for name, defs in VARIANTS.items():
            if name not in only:
                continue
            b.build(force=True, defines=defs, out=os.path.join(OUTDIR, "libm2kdetect_%s.so" % name))
            print("built", name)
    if "--run" in sys.argv:
        wl = sys.argv[sys.argv.index("--workload") + 1] if "--workload" in sys.argv else "C"
        only = sys.argv[sys.argv.index("--only") + 1].split(",") if "--only" in sys.argv else list(VARIANTS)
        for name in only:
            env = dict(os.environ, M2K_LIB=os.path.join(OUTDIR, "libm2kdetect_%s.so" % name))
            p = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--workload", wl, "--steps", "5", "--warmup", "3",
                                "--no-e2e", "--no-cpu"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
            try:
                j = json.loads(p.stdout.strip().splitlines()[-1])
                print("%-14s scan %.3f ms  %.0f GB/s  frac %.3f  step %.3f ms  combine %.3f  dockerfile %.3f" % (
                    name, j["roofline"]["kernel_ms"], j["roofline"]["achieved"], j["roofline"]["frac"], j["ms_per_step"],
                    j["breakdown_ms"]["combine"], j["breakdown_ms"]["dockerfile"]), flush=True)
            except Exception as e:
                print(name, "FAILED", e, p.stderr[-300:], flush=True)


if __name__ == "__main__":
    main()
