"""Writes a packed tree whose bytes are on the host (a synthetic tree from m2k_synth_tree, or any
m2k_tree with `bytes`) to a real directory -- used to time the packer (`m2k_pack_tree`, the one
`filepath.Walk` replacement) on the bench generator's trees and by the script-oracle baseline."""
import ctypes as C
import os

from . import _abi


def materialize(tree, dst: bytes) -> int:
    """Create `dst` and everything below it; returns the content bytes written.  Only regular files and
    directories can be written (the generator makes nothing else)."""
    t = tree.contents
    if not t.bytes and t.content_bytes:
        raise ValueError("the tree's bytes are not on the host")
    names = C.string_at(t.names, t.n_name_bytes)
    noff = t.entry_name_off[0:t.n_entries + 1]
    first = t.dir_first_entry[0:t.n_dirs + 1]
    etype = bytes(bytearray(t.entry_type[0:t.n_entries]))
    child = t.entry_child[0:t.n_entries]
    base_addr = C.addressof(t.bytes.contents) if t.bytes else 0
    dir_path = [b""] * t.n_dirs
    dir_path[0] = dst
    os.makedirs(dst, exist_ok=True)
    written = 0
    for d in range(t.n_dirs):
        base = dir_path[d]
        for e in range(first[d], first[d + 1]):
            p = base + b"/" + names[noff[e]:noff[e + 1]]
            ty = etype[e] & 0xF
            if ty == _abi.T_DIR:
                os.mkdir(p)
                dir_path[child[e]] = p
            elif ty == _abi.T_REG:
                f = child[e]
                ln = t.file_len[f]
                with open(p, "wb") as fh:
                    if ln:
                        fh.write(C.string_at(base_addr + t.file_off[f], ln))
                written += ln
            else:
                raise ValueError("only regular files and directories can be materialised")
    return written
