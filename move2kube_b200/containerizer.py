"""Mirror of internal/containerizer (reference: containerizer.go:37-93, dockerfilecontainerizer.go:40-83,
s2icontainerizer.go:41-84, cnbcontainerizer.go:38-63, reusecontainerizer.go:35-37) over the GPU detection.

Same names, argument meaning and error behaviour as the Go interfaces: `Init(path)` discovers option
directories, `GetTargetOptions(plan, path)` returns the option directories that accept `path`
(no error return: "no options" is an empty list), `GetContainerizationOptions` keeps the fixed
order NewDockerfile, S2I, CNB.  The built-in detect scripts are not executed: their predicates were
evaluated for every directory at once by the kernels.  Detect scripts the USER put in the input
tree are arbitrary shell, so those (and only those) are run on the host exactly as the reference
runs them; CNB needs a container runtime and stays a host pass-through hook.
"""
import logging
import os
import subprocess

from . import _abi, detection, plan as plantypes

log = logging.getLogger("move2kube_b200")

dockerfileDetectScript = "m2kdfdetect.sh"
s2iDetectScript = "m2ks2idetect.sh"

# common.AssetsPath (internal/common/constants.go:70-72): where the built-in option directories live
AssetsPath = os.path.join("m2k-temp", "m2kassets")


def _builtin_index(option_dir: str, kind: str):
    """If option_dir is <assets>/<kind>/<name> for a built-in name, its bit index; else None."""
    names = _abi.DF_NAMES if kind == "dockerfiles" else _abi.S2I_NAMES
    base = os.path.basename(option_dir)
    parent = os.path.basename(os.path.dirname(option_dir))
    if parent == kind and base in names:
        return names.index(base)
    return None


def _find_scripts(path: str, script: str):
    """common.GetFilesByName(path, [script]) for the small assets directory (host walk, Walk order)."""
    out = []
    if not os.path.isdir(path):
        log.warning("Unable to fetch files to recognize detect scripts : %s does not exist", path)
        return out
    for dirpath, dirnames, filenames in os.walk(path):
        dirnames.sort()
        entries = sorted(filenames + [d for d in dirnames if os.path.islink(os.path.join(dirpath, d))])
        for fn in entries:
            if fn == script:
                out.append(os.path.join(dirpath, fn))
    # os.walk yields a directory's files before its sub-directories; filepath.Walk interleaves by name
    out.sort(key=lambda p: p.split(os.sep))
    return out


class _ScriptContainerizer:
    script = None
    kind = None
    feature_bit = None
    opt_shift = 0
    build_type = None

    def __init__(self):
        self.options = []  # option directories, in Init discovery order
        self._det = None

    def GetContainerBuildStrategy(self):
        return self.build_type

    def Init(self, path: str, det: "detection.Detection" = None):
        """dockerfilecontainerizer.go:50-59 -- option dirs = Dir(f) for every detect script under path."""
        if det is not None and os.fsencode(path) == det.broot:
            self._det = det
            for e in det.entries_with(self.feature_bit):
                self.options.append(os.fsdecode(os.path.dirname(det.entry_path(e))))
            return
        for f in _find_scripts(path, self.script):
            self.options.append(os.path.dirname(f))

    def detect(self, scriptDir: str, directory: str, det: "detection.Detection" = None):
        """dockerfilecontainerizer.go:76-83: (stdout, err) of the option's detect script for `directory`.
        For a built-in option the answer comes from the detection pass (m2k_option_json: the script's stdout
        byte for byte, nothing executed) -- this is the call GetContainer makes again at translate time
        (dockerfilecontainerizer.go:105-129, s2icontainerizer.go:97-115).  User-supplied option directories
        are arbitrary shell and are executed exactly as the reference executes them."""
        det = det or self._det
        idx = _builtin_index(scriptDir, self.kind)
        if idx is not None and det is not None:
            d = det.dir_id(directory)
            if d >= 0:
                out = det.option_json(d, self.opt_shift + idx)
                if out is None:
                    return "", RuntimeError("exit status 1")
                return out.decode("utf-8", "replace"), None
        try:
            p = subprocess.run(["/bin/sh", self.script, directory], cwd=scriptDir, stdout=subprocess.PIPE)
        except OSError as e:
            return "", e
        if p.returncode != 0:
            return p.stdout.decode("utf-8", "replace"), RuntimeError("exit status %d" % p.returncode)
        return p.stdout.decode("utf-8", "replace"), None

    def GetTargetOptions(self, plan, path: str, det: "detection.Detection" = None):
        det = det or self._det
        d = det.dir_id(path) if det is not None else -1
        targetOptions = []
        for opt in self.options:
            idx = _builtin_index(opt, self.kind)
            if idx is not None:
                if d < 0:
                    raise RuntimeError("directory %r is not part of the detected tree" % path)
                if det.dir_opts[d] >> (self.opt_shift + idx) & 1:
                    targetOptions.append(opt)
                continue
            output, err = self.detect(opt, path)
            if err is not None:
                log.debug("%s detector cannot containerize %s Error: %s", opt, path, err)
                continue
            targetOptions.append(opt)
        return targetOptions


class DockerfileContainerizer(_ScriptContainerizer):
    script, kind, feature_bit, opt_shift = dockerfileDetectScript, "dockerfiles", _abi.N_USER_DF, 0
    build_type = plantypes.DockerFileContainerBuildTypeValue


class S2IContainerizer(_ScriptContainerizer):
    script, kind, feature_bit, opt_shift = s2iDetectScript, "s2i", _abi.N_USER_S2I, 12
    build_type = plantypes.S2IContainerBuildTypeValue


class CNBContainerizer:
    """cnbcontainerizer.go:41-63.  The detector lives inside third-party builder images and needs a
    container runtime (cnb/provider.go:68-78); `provider` is the host hook, absent = every provider
    errored = no CNB options (what the reference returns on a box without docker/podman/pack/runc)."""
    builders = ["cloudfoundry/cnb:cflinuxfs3", "gcr.io/buildpacks/builder"]
    provider = None  # callable(path, builder) -> bool
    _cache = {}

    def GetContainerBuildStrategy(self):
        return plantypes.CNBContainerBuildTypeValue

    def Init(self, path, det=None):
        pass

    def GetTargetOptions(self, plan, path, det=None):
        if path in CNBContainerizer._cache:
            return CNBContainerizer._cache[path]
        supported = []
        if CNBContainerizer.provider is not None:
            supported = [b for b in self.builders if CNBContainerizer.provider(path, b)]
        CNBContainerizer._cache[path] = supported
        return supported


class ReuseContainerizer:
    """reusecontainerizer.go:35-37: never offers options."""

    def GetContainerBuildStrategy(self):
        return plantypes.ReuseContainerBuildTypeValue

    def Init(self, path, det=None):
        pass

    def GetTargetOptions(self, plan, path, det=None):
        return []


class ContainerizationOption:
    """containerizer.go:50-53"""

    def __init__(self, ContainerizationType, TargetOptions):
        self.ContainerizationType = ContainerizationType
        self.TargetOptions = TargetOptions

    def __repr__(self):
        return "ContainerizationOption(%r, %r)" % (self.ContainerizationType, self.TargetOptions)


class Containerizers:
    """containerizer.go:45-76"""

    def __init__(self):
        self.containerizers = []
        self.det = None
        self._own_det = False

    def close(self):
        if self._own_det and self.det is not None:
            self.det.close()
        self.det = None
        self._own_det = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def InitContainerizers(self, path: str, det: "detection.Detection" = None):
        """containerizer.go:56-62: Init(path) then Init(AssetsPath) for each of the four containerizers."""
        if det is None and os.path.isdir(path) and not os.path.islink(path):
            det = detection.detect(path)  # owned by this object: released by close() / when it is collected
            self._own_det = True
        self.det = det
        self.containerizers = [DockerfileContainerizer(), S2IContainerizer(), CNBContainerizer(), ReuseContainerizer()]
        for c in self.containerizers:
            c.Init(path, det)
            c.Init(AssetsPath, det)

    def has_user_options(self) -> bool:
        for c in self.containerizers:
            if isinstance(c, _ScriptContainerizer):
                if any(_builtin_index(o, c.kind) is None for o in c.options):
                    return True
        return CNBContainerizer.provider is not None

    def GetContainerizationOptions(self, plan, sourcepath: str):
        cops = []
        for c in self.containerizers:
            targetoptions = c.GetTargetOptions(plan, sourcepath, self.det)
            if len(targetoptions) != 0:
                cops.append(ContainerizationOption(c.GetContainerBuildStrategy(), targetoptions))
        return cops
