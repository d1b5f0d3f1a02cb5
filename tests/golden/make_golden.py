#!/usr/bin/env python3
"""Generates tests/golden/*.json.  Run in the build container, where /root/reference exists:

    python oracle/build_ref.py && python tests/golden/make_golden.py

For every fixture tree it materialises the spec in a temp dir, runs the script oracle
(oracle/script_oracle.py: the reference's control flow + the reference's own detect scripts,
both /bin/sh flavours) and stores spec + expected answers, all paths relative to the tree
root so the fixture is machine-independent.

Fixtures and where the reference pins them (paths under /root/reference):
  samples                       samples/ (BASELINE.json configs[0]); SURVEY.md App. B
  samples_nodejs[_pre]          internal/source/any2kube_test.go:97-212, internal/move2kube/planner_test.go:99-128
  ignorecase1 / ignorecase2     internal/source/any2kube_test.go:214-286 + testdata/*/.m2kignore
  ignorecontents                any2kube_test.go:288-332 + testdata/m2kignoreforignorecontents
  multiignores                  any2kube_test.go:334-391 + testdata/testmultiplem2kignores.tar
  emptydir / nonexistent        any2kube_test.go:38-70
  nasty, userscript, randomNNN  ours: edge cases of SURVEY.md App. A and random small trees
  pairwise                      ours: every single marker and every pair of markers, each in its own directory
"""
import json
import os
import shutil
import sys
import tarfile
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(REPO, "oracle"))
sys.path.insert(0, os.path.join(REPO, "tests"))
import gopath  # noqa: E402
import script_oracle as so  # noqa: E402
import treespec  # noqa: E402

REF = "/root/reference"


def rel(p: bytes, root: bytes) -> str:
    return os.fsdecode(os.path.relpath(p, root))


def opt_name(p: bytes, root: bytes) -> str:
    assets = so.ASSETS.encode()
    if p.startswith(assets + b"/"):
        return "m2kassets/" + os.fsdecode(p[len(assets) + 1:])
    return "tree:" + rel(p, root)


def main_candidates(d: bytes):
    """Files a `grep -lRe __main__ D | awk '/.py$/'` can list, excluding anything reached through a
    symlinked directory (the packed tree does not follow those)."""
    out = []
    for dirpath, dirnames, filenames in os.walk(d):
        dirnames.sort()
        for fn in sorted(filenames):
            full = os.path.join(dirpath, fn)
            if not full.endswith(b"py"):
                continue
            try:
                if os.path.isfile(full) and b"__main__" in open(full, "rb").read():
                    out.append(os.fsdecode(os.path.relpath(full, d)))
            except OSError:
                pass
    return out


def has_dir_symlink(root: bytes) -> bool:
    for dirpath, dirnames, filenames in os.walk(root):
        for x in dirnames + filenames:
            full = os.path.join(dirpath, x)
            if os.path.islink(full) and os.path.isdir(full):
                return True
    return False


def run_oracle(root: str, pre=()):
    broot = os.fsencode(root)
    out = {}
    for flavor in ("bash", "dash"):
        trace = []
        pre_abs = [gopath.join(broot, os.fsencode(p)) if p != "." else broot for p in pre]
        services, err, cs, (ign_d, ign_c) = so.get_service_options(broot, pre_containerized=pre_abs,
                                                                  flavor=flavor, trace=trace)
        exp = {
            "error": None if err is None else str(err),
            "ignore_dirs": [rel(p, broot) for p in ign_d],
            "ignore_contents": [rel(p, broot) for p in ign_c],
            "trace": [[rel(p, broot), action, {bt: [opt_name(t, broot) for t in to] for bt, to in cops}]
                      for p, action, cops in trace],
            "services": [[rel(s["sourceArtifacts"]["SourceCode"][0], broot), s["serviceName"],
                          s["containerBuildType"], [opt_name(t, broot) for t in s["targetOptions"]]]
                         for s in services],
            "main_script": {},
            # what every accepting detect script printed on stdout, per service directory and option: the JSON that
            # GetContainer re-runs the script for at translate time (dockerfilecontainerizer.go:105-129, s2icontainerizer.go:97-115)
            "outputs": {},
        }
        dir_links = os.path.isdir(root) and has_dir_symlink(broot)
        for p, action, cops in trace:
            if action != "service":
                continue
            outs = {}
            for c in cs.list:
                if isinstance(c, so.ScriptContainerizer):
                    for (od, d), stdout in c.outputs.items():
                        if d == p:
                            outs[opt_name(od, broot)] = stdout.decode("latin-1")  # bytes kept one to one
            exp["outputs"][rel(p, broot)] = outs
            ms = so.main_script_rel_path(cs, p)
            if ms:
                exp["main_script"][rel(p, broot)] = {
                    "script_said": ms,
                    "candidates": None if dir_links else main_candidates(p),
                }
        out[flavor] = exp
    return out


def fixture(spec, out_dir, tmp):
    root = os.path.join(tmp, spec["name"])
    if os.path.exists(root):
        shutil.rmtree(root)
    if not spec.get("nonexistent"):
        treespec.materialize(spec, root)
    spec["expect"] = run_oracle(root, spec.get("pre", ()))
    with open(os.path.join(out_dir, spec["name"] + ".json"), "w") as f:
        json.dump(spec, f, indent=1, sort_keys=True)
    n = len(spec["expect"]["bash"]["services"])
    print("%-28s entries=%-4d services(bash)=%d" % (spec["name"], len(spec["entries"]), n))


def nasty_spec():
    L = treespec.lit
    e = [
        # a DIRECTORY named x.go two levels down: find -name matches any type -> golang accepts
        {"p": "godir/a/x.go", "t": "d"},
        # a file named exactly ".go" (find has no dotfile rule)
        {"p": "dotgo", "t": "d"}, {"p": "dotgo/.go", "t": "f", "size": 3},
        # ".war" dotfile: the shell glob skips it -> war rejects
        {"p": "dotwar", "t": "d"}, {"p": "dotwar/.war", "t": "f", "size": 3},
        # a.war broken symlink sorts first -> war rejects although z.war is real
        {"p": "warfirstbroken", "t": "d"},
        {"p": "warfirstbroken/a.war", "t": "l", "to": "nonexistent-target"},
        {"p": "warfirstbroken/z.war", "t": "f", "size": 10},
        # first match exists -> accept (three war containerizers)
        {"p": "warok", "t": "d"}, {"p": "warok/app.war", "t": "f", "size": 10},
        {"p": "warok/zz.war", "t": "l", "to": "nonexistent-target"},
        # .war directory counts ([ -e ])
        {"p": "wardir", "t": "d"}, {"p": "wardir/x.war", "t": "d"},
        # package.json -> symlink to a regular file: accepts; -> symlink to a directory: rejects
        {"p": "linkfile", "t": "d"}, {"p": "linkfile/real.json", "t": "f", "size": 5},
        {"p": "linkfile/package.json", "t": "l", "to": "real.json"},
        {"p": "linkdir", "t": "d"}, {"p": "linkdir/sub", "t": "d"},
        {"p": "linkdir/package.json", "t": "l", "to": "sub"},
        # package.json that is a directory: [ -f ] false
        {"p": "dirmarker", "t": "d"}, {"p": "dirmarker/package.json", "t": "d"},
        # case sensitivity of marker names
        {"p": "casemarker", "t": "d"}, {"p": "casemarker/Package.json", "t": "f", "size": 5},
        # gradle beats maven/java for s2i java; ant too
        {"p": "gradlemaven", "t": "d"}, {"p": "gradlemaven/build.gradle", "t": "f", "size": 5},
        {"p": "gradlemaven/pom.xml", "t": "f", "size": 5},
        {"p": "antjava", "t": "d"}, {"p": "antjava/build.xml", "t": "f", "size": 5},
        {"p": "antjava/src/A.java", "t": "f", "size": 5},
        {"p": "purejava", "t": "d"}, {"p": "purejava/deep/er/A.java", "t": "f", "size": 5},
        # go.mod only (no *.go): s2i golang yes, dockerfile golang no
        {"p": "gomodonly", "t": "d"}, {"p": "gomodonly/go.mod", "t": "f", "size": 5},
        # python: each special file; names ending in "py" are awk /.py$/ candidates
        {"p": "py_req", "t": "d"}, {"p": "py_req/requirements.txt", "t": "f", "size": 5},
        {"p": "py_req/zz/copy", "t": "f", "size": 50, "main": True},
        {"p": "py_setup", "t": "d"}, {"p": "py_setup/setup.py", "t": "f", "size": 50, "main": True},
        {"p": "py_env", "t": "d"}, {"p": "py_env/environment.yml", "t": "f", "size": 5},
        {"p": "py_env/main.py", "t": "f", "size": 80},
        {"p": "py_pipfile", "t": "d"}, {"p": "py_pipfile/Pipfile", "t": "f", "size": 5},
        {"p": "py_pipfile/bin.py", "t": "f", "b64": "AAECA19fbWFpbl9fAP8="},
        # Kelvin sign / long s / mixed case in ignore entries (strings.EqualFold)
        {"p": "Kdir", "t": "d"}, {"p": "Kdir/package.json", "t": "f", "size": 5},
        {"p": "sdir", "t": "d"}, {"p": "sdir/package.json", "t": "f", "size": 5},
        {"p": "MiXeD", "t": "d"}, {"p": "MiXeD/package.json", "t": "f", "size": 5},
        {"p": "mixed", "t": "d"}, {"p": "mixed/Gemfile", "t": "f", "size": 5},
        {"p": "keepme", "t": "d"}, {"p": "keepme/sub/package.json", "t": "f", "size": 5},
        {"p": "dotdot", "t": "d"}, {"p": "dotdot/x/pom.xml", "t": "f", "size": 5},
        {"p": "crlf", "t": "d"}, {"p": "crlf/p/package.json", "t": "f", "size": 5},
        {"p": "crlf/q/package.json", "t": "f", "size": 5},
        L("crlf/.m2kignore", b"p/\r\np/*\r\n\xc2\xa0q\xe2\x80\x83\r\n q* \r\n"),
        L("dotdot/y/.m2kignore", b"../x/\n../x/*\n"),
        L(".m2kignore", ("\u212adir/\n\u212adir/*\n\u017fdir\n\u017fDIR*\nmixed\nMIXED/*\nkeepme\n"
                         "nonexistent/path/*\nkeepme/**\n\n").encode()),
    ]
    return {"name": "nasty", "entries": e}


def pairwise_spec():
    """The 18 detect scripts' truth table, pinned beyond the fixtures above: one first-level directory for every single
    marker and every PAIR of markers the scripts look at (own files, files anywhere below, a *.war, a python main), and
    every subset of the four that s2i/java weighs against each other.  The root is ignored but descended (an empty
    .m2kignore line, as in `nasty`), so each directory meets the scripts on its own."""
    markers = {
        "pipfile": [("Pipfile", {})], "buildxml": [("build.xml", {})], "gradle": [("build.gradle", {})], "pom": [("pom.xml", {})],
        "pkgjson": [("package.json", {})], "reqs": [("requirements.txt", {})], "setuppy": [("setup.py", {})], "envyml": [("environment.yml", {})],
        "gemfile": [("Gemfile", {})], "gomod": [("go.mod", {})],
        "go": [("x.go", {})], "godeep": [("sub/deeper/y.go", {})], "php": [("index.php", {})], "phpdeep": [("sub/z.php", {})],
        "java": [("A.java", {})], "javadeep": [("src/main/B.java", {})], "war": [("app.war", {})],
        "mainpy": [("tool/run.py", {"size": 60, "main": True})],
    }
    names = sorted(markers)
    combos = [()] + [(a,) for a in names] + [(a, b) for i, a in enumerate(names) for b in names[i + 1:]]
    four = ("gradle", "buildxml", "pom", "javadeep")
    for mask in range(16):
        c = tuple(sorted(four[k] for k in range(4) if mask >> k & 1))
        if len(c) > 2:
            combos.append(c)
    e = [treespec.lit(".m2kignore", b"\n")]
    for c in combos:
        d = "_".join(c) if c else "none"
        e.append({"p": d, "t": "d"})
        for m in c:
            for rel_path, extra in markers[m]:
                ent = {"p": d + "/" + rel_path, "t": "f", "size": 7}
                ent.update(extra)
                e.append(ent)
    return {"name": "pairwise", "entries": e}


def userscript_spec():
    L = treespec.lit
    script = (b"#!/bin/sh\nif [ -f \"$1/custom.marker\" ]; then\n  echo '{\"port\": 1234}'\n  exit 0\nfi\nexit 1\n")
    e = [
        L("containerizers/mydf/m2kdfdetect.sh", script),
        L("containerizers/mys2i/m2ks2idetect.sh", script),
        {"p": "appA", "t": "d"}, {"p": "appA/custom.marker", "t": "f", "size": 1},
        {"p": "appB", "t": "d"}, {"p": "appB/custom.marker", "t": "f", "size": 1},
        {"p": "appB/package.json", "t": "f", "size": 5},
        {"p": "appC", "t": "d"}, {"p": "appC/readme", "t": "f", "size": 5},
    ]
    return {"name": "userscript", "entries": e}


def reference_plans(out_dir):
    """The reference's own golden YAML plans for this path, as JSON (services only):
    internal/move2kube/testdata/expectedplanfornodejsapp.yaml        <- planner_test.go:99-128
    internal/source/testdata/expectedservicesfor*.yaml               <- any2kube_test.go:97-286"""
    import yaml
    files = {
        "planfornodejsapp": ("internal/move2kube/testdata/expectedplanfornodejsapp.yaml", "samples_nodejs"),
        "servicesfornodejsapp": ("internal/source/testdata/expectedservicesfornodejsapp.yaml", "samples_nodejs"),
        "servicesfornodejsappwithm2kignorecase1": ("internal/source/testdata/expectedservicesfornodejsappwithm2kignorecase1.yaml", "ignorecase1"),
        "servicesforjavamavenappwithm2kignorecase2": ("internal/source/testdata/expectedservicesforjavamavenappwithm2kignorecase2.yaml", "ignorecase2"),
    }
    out = {}
    for key, (rel_path, tree) in files.items():
        y = yaml.safe_load(open(os.path.join(REF, rel_path)))
        # "text": the golden file as the reference's WritePlan stored it (planio's writer is compared with it byte for byte)
        out[key] = {"source": rel_path, "tree": tree, "services": y["spec"]["inputs"]["services"],
                    "text": open(os.path.join(REF, rel_path)).read()}
    with open(os.path.join(out_dir, "_reference_plans.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("reference plans: %d" % len(out))


def main():
    out_dir = HERE
    if len(sys.argv) > 2 and sys.argv[1] == "--only":  # python make_golden.py --only pairwise[,nasty,userscript]
        tmp = tempfile.mkdtemp(prefix="m2kgold-")
        try:
            for name in sys.argv[2].split(","):
                fixture({"pairwise": pairwise_spec, "nasty": nasty_spec, "userscript": userscript_spec}[name](), out_dir, tmp)
        finally:
            shutil.rmtree(tmp, ignore_errors=True)
        return
    reference_plans(out_dir)
    tmp = tempfile.mkdtemp(prefix="m2kgold-")
    try:
        s = treespec.spec_from_disk("samples", os.path.join(REF, "samples"))
        fixture(s, out_dir, tmp)
        s = treespec.spec_from_disk("samples_nodejs", os.path.join(REF, "samples", "nodejs"))
        fixture(s, out_dir, tmp)
        s = treespec.spec_from_disk("samples_nodejs_pre", os.path.join(REF, "samples", "nodejs"))
        s["pre"] = ["."]
        fixture(s, out_dir, tmp)
        td = os.path.join(REF, "internal", "source", "testdata")
        fixture(treespec.spec_from_disk("ignorecase1", os.path.join(td, "nodejsappwithm2kignorecase1")), out_dir, tmp)
        fixture(treespec.spec_from_disk("ignorecase2", os.path.join(td, "javamavenappwithm2kignorecase2")), out_dir, tmp)
        rules = open(os.path.join(td, "m2kignoreforignorecontents"), "rb").read()
        fixture({"name": "ignorecontents", "entries": [
            {"p": "includeme/excludeme", "t": "d"},
            treespec.lit(".m2kignore", rules),
            treespec.lit("includeme/excludeme/package.json", b"this is ' invalid json")]}, out_dir, tmp)
        with tarfile.open(os.path.join(td, "testmultiplem2kignores.tar")) as tf:
            tf.extractall(os.path.join(tmp, "tarx"))
        fixture(treespec.spec_from_disk("multiignores", os.path.join(tmp, "tarx", "testmultiplem2kignores")), out_dir, tmp)
        fixture({"name": "emptydir", "entries": []}, out_dir, tmp)
        fixture({"name": "nonexistent", "entries": [], "nonexistent": True}, out_dir, tmp)
        fixture(nasty_spec(), out_dir, tmp)
        fixture(userscript_spec(), out_dir, tmp)
        fixture(pairwise_spec(), out_dir, tmp)
        for seed in range(1, 41):
            fixture(treespec.random_spec(seed, n_dirs=6 + seed % 9, n_files=14 + (seed * 7) % 30,
                                         ignore_files=seed % 4), out_dir, tmp)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
