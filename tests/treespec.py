"""Tree specs: a small JSON description of a directory tree that tests materialise on disk.

A spec is {"name": str, "entries": [entry...], "pre": [rel dir...]?} where an entry is
  {"p": rel_path, "t": "d"}                          directory
  {"p": rel_path, "t": "f", "b64": ...}              file with literal bytes
  {"p": rel_path, "t": "f", "size": n, "main": bool} file with synthetic bytes (filler lines;
                                                     contains "__main__" iff main)
  {"p": rel_path, "t": "l", "to": target}            symlink
Paths are str with surrogateescape for non-UTF-8 bytes.  Parents are created implicitly.

Reference-derived specs (samples/, internal/source/testdata/*) keep names, types, sizes and
the one content bit the detect scripts look at (`__main__`); `.m2kignore` files, which ARE the
reference's test vectors for the ignore semantics, are kept literally.
"""
import base64
import os
import random


def filler(size: int, has_main: bool, salt: int = 0) -> bytes:
    if size <= 0 and not has_main:
        return b""
    rng = random.Random(size * 1000003 + salt)
    lines = []
    total = 0
    marker = b"if __name__ == '__main__':\n"
    if has_main:
        lines.append(marker)
        total += len(marker)
    words = [b"alpha", b"beta", b"gamma", b"delta", b"return", b"value", b"x = 1", b"//", b"{", b"}"]
    while total < size:
        ln = b" ".join(rng.choice(words) for _ in range(rng.randint(1, 8))) + b"\n"
        lines.append(ln)
        total += len(ln)
    if has_main and len(lines) > 2:
        # move the marker somewhere inside the file
        k = rng.randrange(len(lines))
        lines[0], lines[k] = lines[k], lines[0]
    data = b"".join(lines)
    if not has_main:
        data = data[:size]
    return data


def entry_bytes(e) -> bytes:
    if "b64" in e:
        return base64.b64decode(e["b64"])
    return filler(int(e.get("size", 0)), bool(e.get("main", False)), salt=len(e["p"]))


def materialize(spec, root: str) -> str:
    """Create the tree under `root` (created if missing).  Returns root."""
    os.makedirs(root, exist_ok=True)
    broot = os.fsencode(root)
    for e in spec["entries"]:
        p = os.path.join(broot, os.fsencode(e["p"])) if e["p"] not in ("", ".") else broot
        t = e["t"]
        if t == "d":
            os.makedirs(p, exist_ok=True)
        elif t == "f":
            os.makedirs(os.path.dirname(p), exist_ok=True)
            with open(p, "wb") as f:
                f.write(entry_bytes(e))
            if "mode" in e:
                os.chmod(p, int(e["mode"], 8))
        elif t == "l":
            os.makedirs(os.path.dirname(p), exist_ok=True)
            if os.path.lexists(p):
                os.unlink(p)
            os.symlink(os.fsencode(e["to"]), p)
        else:
            raise ValueError(t)
    return root


def lit(path: str, data: bytes, mode=None):
    e = {"p": path, "t": "f", "b64": base64.b64encode(data).decode()}
    if mode:
        e["mode"] = mode
    return e


def spec_from_disk(name: str, root: str, literal_names=(".m2kignore",)):
    """Describe an on-disk tree (used on reference fixtures by make_golden.py)."""
    entries = []
    broot = os.fsencode(root)
    for dirpath, dirnames, filenames in os.walk(broot):
        dirnames.sort()
        rel = os.path.relpath(dirpath, broot)
        if rel != b".":
            entries.append({"p": os.fsdecode(rel), "t": "d"})
        for fn in sorted(filenames) + [d for d in dirnames if os.path.islink(os.path.join(dirpath, d))]:
            full = os.path.join(dirpath, fn)
            relp = os.fsdecode(os.path.normpath(os.path.join(rel, fn)))
            if os.path.islink(full):
                entries.append({"p": relp, "t": "l", "to": os.fsdecode(os.readlink(full))})
                continue
            data = open(full, "rb").read()
            if os.fsdecode(fn) in literal_names:
                entries.append(lit(relp, data))
            else:
                entries.append({"p": relp, "t": "f", "size": len(data), "main": b"__main__" in data})
    return {"name": name, "entries": entries}


# ---- random small trees ---------------------------------------------------------------

MARKERS = ["Pipfile", "build.xml", "build.gradle", "pom.xml", "package.json", "requirements.txt",
           "setup.py", "environment.yml", "Gemfile", "go.mod"]
PLAIN = ["README.md", "main.py", "app.py", "util.py", "index.js", "main.go", "lib.go", "App.java",
         "index.php", "app.rb", "x.war", "a.war", ".war", ".go", "notes.txt", "Makefile", "data.bin",
         "config.yml", "deploy.yaml", "Dockerfile", "PACKAGE.JSON", "Pom.xml", "x.GO", "y.phps",
         "z.java.bak", "go.mod.bak", "copy", "py"]
DIRS = ["src", "app", "lib", "cmd", "pkg", "web", "api", "x.go", "d.php", "svc", "Src", "APP", "vendor",
        "test", "a", "b", "c"]


def random_spec(seed: int, n_dirs=10, n_files=28, ignore_files=2):
    rng = random.Random(seed)
    dirs = ["."]
    entries = []
    for _ in range(n_dirs):
        parent = rng.choice(dirs)
        nm = rng.choice(DIRS)
        p = os.path.normpath(os.path.join(parent, nm))
        if p in dirs:
            continue
        dirs.append(p)
        entries.append({"p": p, "t": "d"})
    used = set(dirs)
    for _ in range(n_files):
        parent = rng.choice(dirs)
        nm = rng.choice(MARKERS) if rng.random() < 0.35 else rng.choice(PLAIN)
        p = os.path.normpath(os.path.join(parent, nm))
        if p in used:
            continue
        used.add(p)
        r = rng.random()
        if r < 0.08 and len(entries) > 0:
            # symlink: to an existing file or dir of this tree (relative), or dangling
            k = rng.random()
            if k < 0.3:
                tgt = "nonexistent-zzz"
            else:
                tgt = os.path.relpath(rng.choice(entries)["p"], parent)
            entries.append({"p": p, "t": "l", "to": tgt})
        else:
            has_main = nm.endswith(".py") and rng.random() < 0.5
            entries.append({"p": p, "t": "f", "size": rng.randint(0, 400), "main": has_main})
    for _ in range(ignore_files):
        parent = rng.choice(dirs)
        p = os.path.normpath(os.path.join(parent, ".m2kignore"))
        if p in used:
            continue
        used.add(p)
        lines = []
        for _ in range(rng.randint(1, 5)):
            tgt = rng.choice(dirs)
            rel = os.path.relpath(tgt, parent)
            form = rng.random()
            if rng.random() < 0.15:
                rel = rel.swapcase()
            if form < 0.35:
                lines.append(rel + "/")
            elif form < 0.7:
                lines.append(rel + "/*")
            elif form < 0.8:
                lines.append(".")
            elif form < 0.9:
                lines.append("  " + rel + "  ")
            else:
                lines.append("/" + rel + "*")
        entries.append(lit(p, ("\n".join(lines) + ("\n" if rng.random() < 0.7 else "")).encode()))
    return {"name": "random%03d" % seed, "entries": entries}
