#!/usr/bin/env python3
"""Generates tests/golden/_reference_dockerfile_corpus.json: every Dockerfile that ships in the reference tree
(its own build Dockerfile, the three under samples/, the twelve templates under internal/assets/dockerfiles) as
"isDockerFile must be true" vectors, and a spread of other files of the same tree (the 18 detect scripts, Go and
YAML sources, samples' manifests) as "must be false" vectors.  The expectations are BY INSPECTION, not by running
the reference (no Go toolchain here; moby/buildkit v0.7.2 is not in the tree): every true file is comments, then
`FROM image[:tag] [AS name]`, and its ENV lines are all `KEY=value`; every false file's first non-comment line
is not a FROM instruction.  Run in the build container:  python tests/golden/make_dockerfile_corpus.py"""
import base64
import glob
import json
import os

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))

TRUE = ["Dockerfile"] + sorted(glob.glob("samples/**/Dockerfile", root_dir=REF, recursive=True)) + \
    sorted(glob.glob("internal/assets/dockerfiles/*/Dockerfile", root_dir=REF))
FALSE = sorted(glob.glob("internal/assets/dockerfiles/*/m2kdfdetect.sh", root_dir=REF)) + \
    sorted(glob.glob("internal/assets/s2i/*/m2ks2idetect.sh", root_dir=REF)) + \
    ["go.mod", "Makefile", "README.md", "LICENSE", "internal/source/dockerfile2kube.go", "internal/source/any2kube.go",
     "samples/nodejs/package.json", "samples/java-maven/pom.xml", "samples/python/main.py", "samples/golang/main.go",
     "samples/docker-compose/docker-compose.yaml", "samples/.m2kignore", ".github/workflows/build.yml"]


def main():
    out = []
    for want, paths in ((True, TRUE), (False, FALSE)):
        for rel in paths:
            p = os.path.join(REF, rel)
            if not os.path.isfile(p):
                continue
            data = open(p, "rb").read()
            out.append({"path": rel, "is_dockerfile": want, "bytes_b64": base64.b64encode(data).decode()})
    with open(os.path.join(HERE, "_reference_dockerfile_corpus.json"), "w") as f:
        json.dump({"provenance": "files of the reference tree; expectations by inspection (see make_dockerfile_corpus.py)", "files": out}, f, indent=1)
    print("%d vectors (%d true)" % (len(out), sum(1 for x in out if x["is_dockerfile"])))


if __name__ == "__main__":
    main()
