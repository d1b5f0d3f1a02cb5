"""Compare a detection result (oracle or product: same field names) with a golden fixture."""
import glob
import json
import os

from move2kube_b200 import _abi

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_goldens():
    out = []
    for p in sorted(glob.glob(os.path.join(GOLDEN_DIR, "*.json"))):
        if os.path.basename(p).startswith("_"):
            continue
        with open(p) as f:
            out.append(json.load(f))
    return out


def golden_ids():
    return [g["name"] for g in load_goldens()]


def has_user_scripts(spec):
    return any(os.path.basename(e["p"]) in ("m2kdfdetect.sh", "m2ks2idetect.sh") and e["t"] != "d"
               for e in spec["entries"])


class TreeView:
    """Relative paths of a packed tree's directories and files."""

    def __init__(self, tree):
        t = tree.contents if hasattr(tree, "contents") else tree
        self.t = t
        self.n_dirs, self.n_entries, self.n_files = t.n_dirs, t.n_entries, t.n_files
        names = bytes(bytearray(t.names[0:t.n_name_bytes])) if t.n_name_bytes else b""
        offs = t.entry_name_off[0:t.n_entries + 1] if t.n_entries else [0]
        self.entry_name = [names[offs[i]:offs[i + 1]] for i in range(t.n_entries)]
        self.dir_rel = [b"."] * t.n_dirs
        for d in range(1, t.n_dirs):
            par = t.dir_parent[d]
            nm = self.entry_name[t.dir_entry[d]]
            self.dir_rel[d] = nm if par == 0 else self.dir_rel[par] + b"/" + nm
        self.file_rel = []
        for f in range(t.n_files):
            e = t.file_entry[f]
            d = t.entry_dir[e]
            self.file_rel.append(self.entry_name[e] if d == 0 else self.dir_rel[d] + b"/" + self.entry_name[e])

    def file_bytes(self, f):
        t = self.t
        off, ln = t.file_off[f], t.file_len[f]
        return bytes(bytearray(t.bytes[off:off + ln])) if ln else b""


def opts_to_names(mask):
    out = {}
    df = ["m2kassets/dockerfiles/" + n for i, n in enumerate(_abi.DF_NAMES) if mask >> i & 1]
    s2i = ["m2kassets/s2i/" + n for i, n in enumerate(_abi.S2I_NAMES) if mask >> (12 + i) & 1]
    if df:
        out["NewDockerfile"] = df
    if s2i:
        out["S2I"] = s2i
    return out


def check(spec, flavor, view, res, root: bytes):
    """res: struct with n_dirs, dir_state, dir_opts, service_dirs, n_services, ignore_*."""
    exp = spec["expect"][flavor]
    r = res.contents if hasattr(res, "contents") else res
    rel_to_dir = {view.dir_rel[d]: d for d in range(view.n_dirs)}
    want_state = {}
    for rel, action, cops in exp["trace"]:
        want_state[os.fsencode(rel)] = (action, cops)
    for d in range(view.n_dirs):
        rel = view.dir_rel[d]
        got = _abi.DIR_STATE_NAMES[r.dir_state[d]]
        want, cops = want_state.get(rel, ("not_visited", {}))
        assert got == want, "dir %r: state %s, golden %s" % (rel, got, want)
        if want == "service":
            assert opts_to_names(r.dir_opts[d]) == cops, "dir %r options" % rel
        elif want in ("nomatch_descend", "nomatch_prune"):
            assert r.dir_opts[d] == 0, "dir %r should have no options" % rel
    for rel in want_state:
        assert rel in rel_to_dir, "golden visits %r which the tree does not have" % rel
    want_services = []
    for rel, name, bt, topts in exp["services"]:
        if not want_services or want_services[-1] != rel:
            want_services.append(rel)
    got_services = [os.fsdecode(view.dir_rel[r.service_dirs[i]]) for i in range(r.n_services)]
    assert got_services == want_services
    got_id = [os.fsdecode(os.path.relpath(r.ignore_dirs[i], root)) for i in range(r.n_ignore_dirs)]
    got_ic = [os.fsdecode(os.path.relpath(r.ignore_contents[i], root)) for i in range(r.n_ignore_contents)]
    assert got_id == exp["ignore_dirs"]
    assert got_ic == exp["ignore_contents"]
    # a11: the main script must be one the reference's grep could have printed
    for rel, ms in exp["main_script"].items():
        d = rel_to_dir[os.fsencode(rel)]
        f = r.dir_main_file[d]
        cands = ms["candidates"]
        if cands is None:
            continue
        if not cands:
            assert f == _abi.NONE
            assert all(v == "" for v in ms["script_said"].values())
        else:
            assert f != _abi.NONE
            got = os.fsdecode(os.path.relpath(view.file_rel[f], os.fsencode(rel)))
            assert got in cands
            assert all(v in cands for v in ms["script_said"].values())


def check_outputs(spec, flavor, view, tree, res):
    """f3: m2k_option_json against what the reference's own scripts printed (expect[flavor]["outputs"], recorded by
    make_golden.py): byte for byte for every option that prints a constant or the war file; for the python options
    the path must be one the reference's grep could have printed (same rule as the main_script check above).
    Returns the number of (directory, option) outputs compared."""
    import json
    from move2kube_b200 import lib
    exp = spec["expect"][flavor]
    rel_to_dir = {view.dir_rel[d]: d for d in range(view.n_dirs)}
    names = ["m2kassets/dockerfiles/" + n for n in _abi.DF_NAMES] + ["m2kassets/s2i/" + n for n in _abi.S2I_NAMES]
    n = 0
    for rel, outs in exp.get("outputs", {}).items():
        d = rel_to_dir[os.fsencode(rel)]
        cands = (exp["main_script"].get(rel) or {}).get("candidates")
        for opt, name in enumerate(names):
            got = lib.option_json(tree, res, d, opt)
            if name not in outs:
                assert got is None, "dir %r option %s: the script did not accept, m2k_option_json answers %r" % (rel, name, got)
                continue
            want = outs[name].encode("latin-1")
            assert got is not None, "dir %r option %s" % (rel, name)
            n += 1
            if name.endswith("/python"):
                key = "main_script_rel_path" if "dockerfiles" in name else "app_file"
                gj, wj = json.loads(got), json.loads(want)
                if cands is None:      # symlinked directories in the tree: the packed tree does not follow them
                    gj.pop(key), wj.pop(key)
                elif cands:
                    assert gj.pop(key) in cands and wj.pop(key) in cands
                assert gj == wj, "dir %r option %s: %r vs %r" % (rel, name, got, want)
                # and the same bytes around the path: key order, spacing, no trailing newline (printf)
                tail = b'", "app_name": "app", "port": 8080}'
                assert got.endswith(tail) and want.endswith(tail) and got.split(b'"')[0:3] == want.split(b'"')[0:3]
            else:
                assert got == want, "dir %r option %s: %r vs %r" % (rel, name, got, want)
    return n


def reference_plans():
    with open(os.path.join(GOLDEN_DIR, "_reference_plans.json")) as f:
        return json.load(f)
