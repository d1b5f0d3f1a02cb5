#!/usr/bin/env python3
"""TEST INFRASTRUCTURE (oracle/) -- never imported by the product path.

Recipe that unpacks the reference's embedded detection assets, the way the reference
itself does at start-up (`CreateAssetsData` -> `UnTarString(assets.Tar, assetsPath)`,
/root/reference/internal/common/utils.go:550-582 and tar.go:75-116), from the base64
tar constant in /root/reference/internal/assets/constants.go:23.

Output goes ONLY to oracle/_ref/m2kassets/ (git-ignored, travels to the GPU box with
gpurun).  Nothing from the reference is committed: this script reads it where it lies.
The 18 detect scripts it unpacks are the reference's own CPU code for the path (the
decision logic is shell, not Go -- SURVEY.md section 0), and `script_oracle.py` executes
them verbatim.

The reference writes the tar without a trailer (tar.go:32-36 defers Close after the
bytes were already taken), so we walk the 512-byte records by hand instead of
trusting `tarfile` to find an end-of-archive marker.
"""
import base64
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("M2K_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref", "m2kassets")


def _tar_entries(blob):
    off = 0
    n = len(blob)
    while off + 512 <= n:
        hdr = blob[off:off + 512]
        if hdr == b"\0" * 512:
            break
        name = hdr[0:100].split(b"\0", 1)[0].decode()
        prefix = hdr[345:500].split(b"\0", 1)[0].decode()
        if prefix:
            name = prefix + "/" + name
        mode = int(hdr[100:108].split(b"\0", 1)[0].strip() or b"0", 8)
        size = int(hdr[124:136].split(b"\0", 1)[0].strip() or b"0", 8)
        typ = hdr[156:157]
        off += 512
        data = blob[off:off + size]
        off += (size + 511) // 512 * 512
        yield name, mode, typ, data


def build(out=OUT, ref=REF):
    src = os.path.join(ref, "internal", "assets", "constants.go")
    if not os.path.exists(src):
        return False
    text = open(src, "r", encoding="utf-8").read()
    m = re.search(r'Tar\s*=\s*`([^`]*)`|Tar\s*=\s*"([^"]*)"', text)
    if not m:
        raise RuntimeError("assets.Tar constant not found in " + src)
    blob = base64.b64decode(m.group(1) or m.group(2))
    n_files = 0
    for name, mode, typ, data in _tar_entries(blob):
        path = os.path.normpath(os.path.join(out, name))
        if typ == b"5" or name.endswith("/"):
            os.makedirs(path, exist_ok=True)
            continue
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "wb") as f:
            f.write(data)
        os.chmod(path, (mode & 0o777) or 0o644)
        n_files += 1
    # sanity: byte-identical to the on-disk copies the tar was generated from
    chk = 0
    for root, _, files in os.walk(out):
        for fn in files:
            rel = os.path.relpath(os.path.join(root, fn), out)
            disk = os.path.join(ref, "internal", "assets", rel)
            if os.path.exists(disk):
                if open(disk, "rb").read() != open(os.path.join(root, fn), "rb").read():
                    raise RuntimeError("unpacked asset differs from reference tree: " + rel)
                chk += 1
    print("oracle/_ref/m2kassets: %d files unpacked, %d verified against %s/internal/assets"
          % (n_files, chk, ref))
    return True


if __name__ == "__main__":
    ok = build()
    if not ok:
        print("reference not present at %s; keeping existing oracle/_ref" % REF)
    sys.exit(0)
