#!/usr/bin/env python3
"""TEST INFRASTRUCTURE (oracle/) -- never imported by the product path.

Script oracle: the reference's Go control flow for source detection restated in Python,
executing the reference's OWN detect scripts (unpacked by build_ref.py into
oracle/_ref/m2kassets) for every decision.  It is the closest thing to running the
reference that this image allows (no Go toolchain), and it is what the golden fixtures
under tests/golden/ were produced with (tests/golden/make_golden.py).

Restated functions (file:line under /root/reference):
  get_files_by_name        internal/common/utils.go:85-120
  get_ignore_paths         internal/source/any2kube.go:151-181
  is_string_present        internal/common/utils.go:308-315
  Containerizers           internal/containerizer/containerizer.go:56-76
  DockerfileContainerizer  internal/containerizer/dockerfilecontainerizer.go:50-83
  S2IContainerizer         internal/containerizer/s2icontainerizer.go:51-84
  CNBContainerizer         internal/containerizer/cnbcontainerizer.go:41-63 (no runtime here -> [])
  get_service_options      internal/source/any2kube.go:43-105
  new_service              internal/source/any2kube.go:143-149, types/plan/plan.go:217-230

Conventions: paths are bytes.  /bin/sh flavour: "bash" (canonical: the project's own UBI8
image has bash as sh) runs the scripts with bash invoked under the name `sh`; "dash" uses
dash.  LC_ALL=C pins the glob collation of `"$1/"*.war`.

[external] filepath.Walk (Go 1.15): Lstat, no symlink following, children in byte order;
a directory whose ReadDir fails is reported to the callback ONCE, with the error (the
reference callback then warns and moves on without detecting it).
"""
import json
import os
import stat
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import gopath  # noqa: E402

ASSETS = os.path.join(HERE, "_ref", "m2kassets")
IGNORE_FILENAME = b".m2kignore"
DF_SCRIPT = b"m2kdfdetect.sh"
S2I_SCRIPT = b"m2ks2idetect.sh"
SKIP_DIR = "SkipDir"


class WalkError(Exception):
    pass


def walk(root: bytes, fn):
    """path/filepath.Walk.  fn(path, st_or_None, err_or_None) -> None | SKIP_DIR | Exception."""
    try:
        st = os.lstat(root)
    except OSError as e:
        r = fn(root, None, e)
        if r is SKIP_DIR or r is None:
            return None
        return r
    r = _walk(root, st, fn)
    if r is SKIP_DIR:
        return None
    return r


def _walk(path, st, fn):
    if not stat.S_ISDIR(st.st_mode):
        return fn(path, st, None)
    err = None
    names = []
    try:
        names = sorted(os.listdir(path))  # bytes in, bytes out; byte order == sort.Strings
    except OSError as e:
        err = e
    r1 = fn(path, st, err)
    if err is not None or r1 is not None:
        return r1
    for name in names:
        child = gopath.join(path, name)
        try:
            cst = os.lstat(child)
        except OSError as e:
            r = fn(child, None, e)
            if r is not None and r is not SKIP_DIR:
                return r
            continue
        r = _walk(child, cst, fn)
        if r is not None:
            if not stat.S_ISDIR(cst.st_mode) or r is not SKIP_DIR:
                return r
    return None


def get_files_by_name(input_path: bytes, names):
    files = []
    try:
        os.stat(input_path)
    except FileNotFoundError as e:
        return files, e
    except OSError:
        pass

    def cb(path, st, err):
        if err is not None and path == input_path:
            return WalkError(str(err))
        if err is not None:
            return None
        if stat.S_ISDIR(st.st_mode):
            return None
        b = gopath.base(path)
        for nm in names:
            if b == nm:
                files.append(path)
        return None

    r = walk(input_path, cb)
    if isinstance(r, Exception):
        return files, r
    return files, None


def get_ignore_paths(input_path: bytes):
    ignore_dirs, ignore_contents = [], []
    file_paths, err = get_files_by_name(input_path, [IGNORE_FILENAME])
    if err is not None:
        return ignore_dirs, ignore_contents
    for fp in file_paths:
        try:
            with open(fp, "rb") as f:
                data = f.read()
        except OSError:
            continue
        gen = gopath.scan_lines(data)
        for raw in gen:
            line = gopath.trim_space(raw)
            d = gopath.dirname(fp)
            if line.endswith(b"*"):
                line = line[:-1]
                ignore_contents.append(gopath.join(d, line))
            else:
                ignore_dirs.append(gopath.join(d, line))
    return ignore_dirs, ignore_contents


def is_string_present(lst, value: bytes) -> bool:
    for v in lst:
        if gopath.equal_fold(v, value):
            return True
    return False


class ScriptContainerizer:
    """Dockerfile / S2I containerizer: option dirs found by name, detect = exec the script."""

    def __init__(self, script_name: bytes, build_type: str, flavor: str):
        self.script = script_name
        self.build_type = build_type
        self.flavor = flavor
        self.option_dirs = []
        self.outputs = {}  # (option_dir, directory) -> stdout of an accepting script

    def init(self, path: bytes):
        files, _ = get_files_by_name(path, [self.script])
        for f in files:
            self.option_dirs.append(gopath.dirname(f))

    def detect(self, script_dir: bytes, directory: bytes):
        if self.flavor == "bash":
            exe, argv0 = "/bin/bash", "sh"
        elif self.flavor == "dash":
            exe, argv0 = "/bin/dash", "sh"
        else:
            raise ValueError(self.flavor)
        env = dict(os.environ)
        env["LC_ALL"] = "C"
        p = subprocess.run([argv0.encode(), self.script, directory], executable=exe,
                           cwd=script_dir, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                           env=env)
        return p.stdout, p.returncode == 0

    def get_target_options(self, directory: bytes):
        opts = []
        for od in self.option_dirs:
            out, ok = self.detect(od, directory)
            if not ok:
                continue
            self.outputs[(od, directory)] = out
            opts.append(od)
        return opts


class CNBContainerizer:
    """cnbcontainerizer.go:41-63 -- every provider errors without a container runtime, so
    IsBuilderSupported is false for both builders: no CNB options on this box."""
    build_type = "CNB"

    def init(self, path):
        pass

    def get_target_options(self, directory):
        return []


class ReuseContainerizer:
    build_type = "Reuse"

    def init(self, path):
        pass

    def get_target_options(self, directory):
        return []


class Containerizers:
    def __init__(self, assets_path: bytes, flavor="bash"):
        self.assets_path = assets_path
        self.flavor = flavor
        self.list = []

    def init_containerizers(self, path: bytes):
        self.list = [ScriptContainerizer(DF_SCRIPT, "NewDockerfile", self.flavor),
                     ScriptContainerizer(S2I_SCRIPT, "S2I", self.flavor),
                     CNBContainerizer(), ReuseContainerizer()]
        for c in self.list:
            c.init(path)
            c.init(self.assets_path)

    def get_containerization_options(self, directory: bytes):
        cops = []
        for c in self.list:
            t = c.get_target_options(directory)
            if len(t) != 0:
                cops.append((c.build_type, t))
        return cops


def new_service(name: bytes, build_type: str, target_options, path: bytes):
    n = name.decode("utf-8", "surrogateescape")
    return {
        "serviceName": n,
        "image": n + ":latest",
        "translationType": "Any2Kube",
        "containerBuildType": build_type,
        "sourceType": ["Directory"],
        "targetOptions": list(target_options),
        "sourceArtifacts": {"SourceCode": [path]},
        "buildArtifacts": {"SourceCode": [path]},
        "updateContainerBuildPipeline": True,
        "updateDeployPipeline": True,
        "serviceRelPath": "/" + n,
    }


def get_service_options(input_path: bytes, pre_containerized=(), assets_path=None, flavor="bash",
                        trace=None):
    """Any2KubeTranslator.GetServiceOptions.  Returns (services, err, containerizers)."""
    if assets_path is None:
        assets_path = ASSETS.encode()
    services = []
    cs = Containerizers(assets_path, flavor)
    cs.init_containerizers(input_path)
    pre = list(pre_containerized)
    ignore_dirs, ignore_contents = get_ignore_paths(input_path)

    def cb(path, st, err):
        if err is not None:
            if trace is not None:
                trace.append((path, "walk_error", []))
            return None
        if not stat.S_ISDIR(st.st_mode):
            return None
        if is_string_present(pre, path):
            if trace is not None:
                trace.append((path, "precontainerized_prune", []))
            return SKIP_DIR
        if is_string_present(ignore_dirs, path):
            if is_string_present(ignore_contents, path):
                if trace is not None:
                    trace.append((path, "ignored_prune", []))
                return SKIP_DIR
            if trace is not None:
                trace.append((path, "ignored_descend", []))
            return None
        cops = cs.get_containerization_options(path)
        if len(cops) == 0:
            if is_string_present(ignore_contents, path):
                if trace is not None:
                    trace.append((path, "nomatch_prune", []))
                return SKIP_DIR
            if trace is not None:
                trace.append((path, "nomatch_descend", []))
            return None
        for bt, topts in cops:
            services.append(new_service(gopath.base(path), bt, topts, path))
        if trace is not None:
            trace.append((path, "service", cops))
        return SKIP_DIR

    r = walk(input_path, cb)
    err = r if isinstance(r, Exception) else None
    return services, err, cs, (ignore_dirs, ignore_contents)


def main_script_rel_path(cs: Containerizers, directory: bytes):
    """The `main_script_rel_path` / `app_file` the python scripts printed for `directory`."""
    out = {}
    for c in cs.list:
        if not isinstance(c, ScriptContainerizer):
            continue
        for (od, d), stdout in c.outputs.items():
            if d != directory:
                continue
            try:
                j = json.loads(stdout.decode("utf-8", "replace"))
            except ValueError:
                continue
            for k in ("main_script_rel_path", "app_file"):
                if k in j:
                    out[c.build_type] = j[k]
    return out


if __name__ == "__main__":
    root = os.fsencode(os.path.abspath(sys.argv[1]))
    tr = []
    svcs, e, cs_, ign = get_service_options(root, trace=tr, flavor=os.environ.get("M2K_SH", "bash"))
    for p, action, cops in tr:
        print(os.fsdecode(p), action, [(bt, [os.fsdecode(gopath.base(t)) for t in to]) for bt, to in cops])
    print("services:", len(svcs), "err:", e)
