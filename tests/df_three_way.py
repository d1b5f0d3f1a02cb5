#!/usr/bin/env python3
"""TEST INFRASTRUCTURE (it runs the oracle): isDockerFile three ways, and the road to pinning it on a machine that has Go.

The parser behind isDockerFile (moby/buildkit v0.7.2) is not in the reference tree and the reference has no test for the
function, so its parity is "unpinned": the oracle (C), the kernel source (CUDA, here built for the host) and
oracle/dockerfile_re.py (Python, on a regex engine) are three derivations by one author.  This tool

  check FILE...            prints the three answers for each file (exit 1 when they differ)
  corpus DIR [--n N]       writes the hand-derived vectors and N grammar-generated files into DIR as 000000.df ...,
                           plus DIR/expected.jsonl ({"file": ..., "is_dockerfile": ...} per line, the three agreeing)
  compare DIR ANSWERS      compares DIR/expected.jsonl with ANSWERS, a file of the same shape written by the REFERENCE:
                           INTEGRATION.md ("Pinning isDockerFile") has the 20-line Go test that produces it from DIR.

Run `corpus` here, carry DIR to a machine with Go and the reference checked out, run the Go test there, bring the
answers back and run `compare`: zero differences pins a12 / f2; any difference names the file to look at.
No GPU is involved: the kernel source is compiled by g++ (tests/host_df/df_host.cc).
"""
import argparse
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))  # this directory
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, ROOT)


def load():
    import dockerfile_re
    import oracle_lib
    import test_df_host_fuzz as fz
    host = fz._build_dfhost("libdfhost.so")

    def three(data: bytes):
        return (bool(oracle_lib.dockerfile_flags(data) & 8), bool(host.dfhost_flags(data, len(data), 0) & 8), dockerfile_re.is_dockerfile(data))
    return three, fz


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    sub = ap.add_subparsers(dest="cmd", required=True)
    p = sub.add_parser("check")
    p.add_argument("files", nargs="+")
    p = sub.add_parser("corpus")
    p.add_argument("dir")
    p.add_argument("--n", type=int, default=20000)
    p.add_argument("--seed", type=int, default=2020)
    p = sub.add_parser("compare")
    p.add_argument("dir")
    p.add_argument("answers")
    a = ap.parse_args()
    if a.cmd == "compare":
        want = {}
        for line in open(os.path.join(a.dir, "expected.jsonl")):
            r = json.loads(line)
            want[r["file"]] = r["is_dockerfile"]
        diff = seen = 0
        for line in open(a.answers):
            r = json.loads(line)
            seen += 1
            if want.get(r["file"]) != r["is_dockerfile"]:
                diff += 1
                print("DIFFERENT %s: reference %s, here %s" % (r["file"], r["is_dockerfile"], want.get(r["file"])))
        print("%d answers, %d expected, %d different" % (seen, len(want), diff))
        return 1 if diff or seen != len(want) else 0
    three, fz = load()
    if a.cmd == "check":
        rc = 0
        for path in a.files:
            data = open(path, "rb").read()
            o, k, p3 = three(data)
            print("%s: oracle %s, kernel source %s, regex-engine derivation %s%s" % (path, o, k, p3, "" if o == k == p3 else "   <-- DIFFER"))
            rc |= not (o == k == p3)
        return rc
    from dockerfile_kats import KATS
    os.makedirs(a.dir, exist_ok=True)
    r = random.Random(a.seed)
    files = [d for _, d, _ in KATS]
    while len(files) < len(KATS) + a.n:
        files.append(fz.gen_file(r) if r.random() < 0.7 else b"FROM a\n" + b"".join(fz.gen_json_line(r) for _ in range(r.randrange(1, 3))))
    with open(os.path.join(a.dir, "expected.jsonl"), "w") as out:
        for i, data in enumerate(files):
            o, k, p3 = three(data)
            if not (o == k == p3):
                print("the three derivations differ on file %06d: oracle %s, kernel source %s, regex-engine %s" % (i, o, k, p3))
                return 1
            name = "%06d.df" % i
            with open(os.path.join(a.dir, name), "wb") as f:
                f.write(data)
            out.write(json.dumps({"file": name, "is_dockerfile": o}) + "\n")
    print("%d files and expected.jsonl written to %s" % (len(files), a.dir))
    return 0


if __name__ == "__main__":
    sys.exit(main())
