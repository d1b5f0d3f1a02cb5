"""TEST INFRASTRUCTURE / MEASUREMENT ONLY -- never imported by the product.

The CPU baseline SURVEY.md section 8(d)(2) asks for: the reference's control flow executing the
reference's OWN detect scripts (oracle/script_oracle.py, bash as /bin/sh, one thread as the
reference is sequential) on a MATERIALISED synthetic tree from the bench generator, timed, and
compared with the plain-C oracle (oracle/m2k_oracle.c) on the same directory -- which pins the C
oracle, and through it the GPU path, against the real scripts at bench-config scale.

    python oracle/script_baseline.py --workload B            # 10 k files / 1 GB on tmpfs
    python oracle/script_baseline.py --files 400 --dirs 60 --bytes 4000000

Needs oracle/_ref/m2kassets (oracle/build_ref.py, from the reference's assets tar).
"""
import argparse
import ctypes as C
import json
import os
import shutil
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, REPO)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(REPO, "tests"))


def materialize(tp, dst: bytes):
    """Write a packed synthetic tree (host bytes present) to the directory `dst`."""
    from move2kube_b200 import synthfs
    return synthfs.materialize(tp, dst)


def script_pass(root: bytes, flavor="bash"):
    import script_oracle as so
    trace = []
    t0 = time.perf_counter()
    services, err, cs, (ign_d, ign_c) = so.get_service_options(root, flavor=flavor, trace=trace)
    dt = time.perf_counter() - t0
    return dt, services, err, trace, ign_d, ign_c


def c_oracle_pass(root: bytes):
    import oracle_lib
    from move2kube_b200 import _abi
    t0 = time.perf_counter()
    rc, tree = oracle_lib.pack_fs(root)
    assert rc == 0
    t1 = time.perf_counter()
    res = oracle_lib.detect(tree, flavor=_abi.SH_BASH, n_threads=1)
    t2 = time.perf_counter()
    return (t1 - t0, t2 - t1), tree, res


# script_oracle trace actions -> M2K_DIR_* (include/m2kdetect.h)
ACTION_STATE = {"precontainerized_prune": 1, "ignored_prune": 2, "ignored_descend": 3, "nomatch_prune": 4,
                "nomatch_descend": 5, "service": 6, "walk_error": 7}


def compare(root: bytes, trace, services, ign_d, ign_c, tree, res):
    """Script oracle vs C oracle: visited directories and their states, option lists, ignore lists."""
    import golden_check as gc
    import oracle_lib
    t, r = tree.contents, res.contents
    path_of = {oracle_lib.dir_path(tree, d): d for d in range(t.n_dirs)}
    seen = {}
    for p, action, cops in trace:
        seen[path_of[p]] = (ACTION_STATE[action], cops)
    bad = []
    for d in range(t.n_dirs):
        st = r.dir_state[d]
        if d in seen:
            want, cops = seen[d]
            if st != want:
                bad.append((d, "state", st, want))
            if want == 6:
                got = gc.opts_to_names(r.dir_opts[d])
                exp = {}
                for bt, topts in cops:
                    if bt == "CNB":
                        continue
                    exp[bt] = ["m2kassets/" + os.fsdecode(os.path.relpath(o, _assets())) for o in topts]
                if got != exp:
                    bad.append((d, "opts", got, exp))
        elif st != 0:
            bad.append((d, "state", st, 0))
    lists_ok = ([r.ignore_dirs[i] for i in range(r.n_ignore_dirs)] == list(ign_d) and
                [r.ignore_contents[i] for i in range(r.n_ignore_contents)] == list(ign_c))
    return bad, lists_ok


def _assets():
    import script_oracle as so
    return so.ASSETS.encode()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default=None, help="bench workload letter (B); overrides the sizes below")
    ap.add_argument("--files", type=int, default=400)
    ap.add_argument("--dirs", type=int, default=60)
    ap.add_argument("--bytes", type=int, default=4_000_000)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--tmp", default="/dev/shm" if os.path.isdir("/dev/shm") else None)
    ap.add_argument("--keep", action="store_true")
    a = ap.parse_args()
    from move2kube_b200 import _abi, lib
    if a.workload:
        import bench
        p = bench.synth_params(a.workload)
    else:
        p = _abi.SynthParams(seed=a.seed, n_files=a.files, n_dirs=a.dirs, max_depth=6, n_ignore_files=max(1, a.dirs // 30),
                             total_bytes=a.bytes)
    tp = lib.synth_tree_host(p, host_bytes=True)
    work = tempfile.mkdtemp(prefix="m2k-synth-", dir=a.tmp)
    root = os.fsencode(os.path.join(work, "tree"))
    try:
        n_bytes = materialize(tp, root)
        t = tp.contents
        nf, nd = t.n_files, t.n_dirs
        lib.tree_free_host(tp)
        s_dt, services, err, trace, ign_d, ign_c = script_pass(root)
        (c_pack, c_detect), tree, res = c_oracle_pass(root)
        bad, lists_ok = compare(root, trace, services, ign_d, ign_c, tree, res)
        out = {
            "tree": {"files": nf, "dirs": nd, "content_bytes": n_bytes, "on": a.tmp or "tmp"},
            "script_oracle": {"seconds": s_dt, "GBps": n_bytes / 1e9 / s_dt, "files_per_s": nf / s_dt,
                              "dirs_visited": len(trace), "services": len(services), "threads": 1, "sh": "bash as sh"},
            "c_oracle": {"pack_seconds": c_pack, "detect_seconds": c_detect, "GBps": n_bytes / 1e9 / (c_pack + c_detect),
                         "threads": 1},
            "agree": {"directory_states_and_options": not bad, "ignore_lists": lists_ok, "mismatches": bad[:10]},
        }
        print(json.dumps(out))
        return 0 if (not bad and lists_ok) else 1
    finally:
        if not a.keep:
            shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    sys.exit(main())
