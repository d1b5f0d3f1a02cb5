"""The plan file (`m2k.plan`): the reference's checkpoint between `plan` and `translate`, and the format
in which detection results from two machines can be compared.

Reference: types/plan/planutils.go:164-201 (ReadPlan / WritePlan), :46-128 (the `m2kpath` struct-tag walk),
:203-268 (IsAssetsPath, GetRelativePath, GetAbsolutePath); types/plan/plan.go:133-215 (yaml keys and their
order); internal/common/utils.go:159-177, 210-251 (WriteYaml with 2-space indent, ReadMove2KubeYaml's
apiVersion check); types/types.go:24-52.

Host-side glue, no GPU: it exists so that a plan written here can be diffed against one written by
`move2kube plan` on a machine that has Go and Docker (tools/plan_diff.py) -- the only road to pinning
the CNB options and isDockerFile against the real reference.
"""
import json
import os
import re

import yaml

from . import containerizer, plan as plantypes

GroupName = "move2kube.konveyor.io"  # types/types.go:29
AssetsDir = "m2kassets"              # internal/common/constants.go:35

_PATH_BUILD_TYPES = ("NewDockerfile", "ReuseDockerfile", "S2I")  # plan.go:209
_PATH_SOURCE_KEYS = ("Kubernetes", "Knative", "DockerCompose", "CfManifest", "CfRunningManifest", "SourceCode", "Dockerfile")  # plan.go:210


def TempPath() -> str:
    """common.TempPath: the directory the assets were unpacked under (AssetsPath = TempPath/m2kassets)."""
    return os.path.dirname(containerizer.AssetsPath)


def IsAssetsPath(path: str) -> bool:
    """planutils.go:203-210"""
    if os.path.isabs(path):
        return path.startswith(TempPath())
    parts = path.split(os.sep)
    return len(parts) > 0 and parts[0] == AssetsDir


def _rel(base: str, target: str) -> str:
    """filepath.Rel for two cleaned paths of the same kind."""
    return os.path.relpath(target, base)


def GetRelativePath(plan, absPath: str) -> str:
    """planutils.go:239-252"""
    if absPath == "" or not os.path.isabs(absPath):
        return absPath
    if IsAssetsPath(absPath):
        return _rel(TempPath(), absPath)
    return _rel(plan.Spec.Inputs.RootDir, absPath)


def GetAbsolutePath(plan, relPath: str) -> str:
    """planutils.go:254-268"""
    if relPath == "" or os.path.isabs(relPath):
        return relPath
    if IsAssetsPath(relPath):
        return os.path.normpath(os.path.join(TempPath(), relPath))
    return os.path.normpath(os.path.join(plan.Spec.Inputs.RootDir, relPath))


def _convert_paths(plan, convert):
    """The `m2kpath` walk of planutils.go:46-128 over the fields that carry the tag (plan.go:164,187-214)."""
    inp = plan.Spec.Inputs
    inp.K8sFiles = [convert(p) for p in inp.K8sFiles]
    inp.QACaches = [convert(p) for p in inp.QACaches]
    inp.TargetInfoArtifacts = {k: [convert(p) for p in v] for k, v in inp.TargetInfoArtifacts.items()}
    tc = plan.Spec.Outputs.Kubernetes.TargetCluster
    tc.Path = convert(tc.Path)
    for services in inp.Services.values():
        for s in services:
            if s.ContainerBuildType in _PATH_BUILD_TYPES:  # `if:ContainerBuildType:in:...`; CNB options are builder names
                s.ContainerizationTargetOptions = [convert(p) for p in s.ContainerizationTargetOptions]
            s.SourceArtifacts = {k: ([convert(p) for p in v] if k in _PATH_SOURCE_KEYS else list(v)) for k, v in s.SourceArtifacts.items()}
            s.BuildArtifacts = {k: [convert(p) for p in v] for k, v in s.BuildArtifacts.items()}
            s.RepoInfo.GitRepoDir = convert(s.RepoInfo.GitRepoDir)
            s.RepoInfo.TargetPath = convert(s.RepoInfo.TargetPath)


def _copy_plan(plan):
    import copy
    return copy.deepcopy(plan)


# ---- encoding: yaml.v3 style (2-space indent, sequences indented under their key, map keys sorted) ----
_PLAIN = re.compile(r"^[A-Za-z0-9_./@$%^&()+=~;<>?-][A-Za-z0-9_./@$%^&()+=~;<>? :,-]*$")


def _scalar(v) -> str:
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, int):
        return str(v)
    s = str(v)
    plain = bool(_PLAIN.match(s)) and not s.endswith((" ", ":")) and ": " not in s and " #" not in s and s[0] not in "-?=~<>&%@"
    if plain:
        try:
            plain = yaml.safe_load(s) == s  # would read back as something else (a number, a bool, null): quote it
        except yaml.YAMLError:
            plain = False
    return s if plain else json.dumps(s, ensure_ascii=False)


def _emit(node, indent, out):
    pad = " " * indent
    if isinstance(node, dict):
        for k, v in node.items():
            if isinstance(v, dict) and v:
                out.append("%s%s:" % (pad, _scalar(k)))
                _emit(v, indent + 2, out)
            elif isinstance(v, list) and v:
                out.append("%s%s:" % (pad, _scalar(k)))
                _emit(v, indent + 2, out)
            else:
                out.append("%s%s: %s" % (pad, _scalar(k), "{}" if isinstance(v, dict) else "[]" if isinstance(v, list) else _scalar(v)))
    else:  # list
        for v in node:
            if isinstance(v, (dict, list)) and v:
                sub = []
                _emit(v, indent + 2, sub)
                out.append("%s- %s" % (pad, sub[0][indent + 2:]))
                out.extend(sub[1:])
            else:
                out.append("%s- %s" % (pad, "{}" if isinstance(v, dict) else "[]" if isinstance(v, list) else _scalar(v)))


def _service_node(s) -> dict:
    """Keys in struct-field order with yaml.v3's omitempty (plan.go:202-215)."""
    n = {"serviceName": s.ServiceName}
    if s.ServiceRelPath:
        n["serviceRelPath"] = s.ServiceRelPath
    n["image"] = s.Image
    n["translationType"] = s.TranslationType
    n["containerBuildType"] = s.ContainerBuildType
    n["sourceType"] = list(s.SourceTypes)
    if s.ContainerizationTargetOptions:
        n["targetOptions"] = list(s.ContainerizationTargetOptions)
    n["sourceArtifacts"] = {k: list(s.SourceArtifacts[k]) for k in sorted(s.SourceArtifacts)}
    if s.BuildArtifacts:
        n["buildArtifacts"] = {k: list(s.BuildArtifacts[k]) for k in sorted(s.BuildArtifacts)}
    n["updateContainerBuildPipeline"] = bool(s.UpdateContainerBuildPipeline)
    n["updateDeployPipeline"] = bool(s.UpdateDeployPipeline)
    r = s.RepoInfo
    if r.GitRepoDir or r.GitRepoURL or r.GitRepoBranch or r.TargetPath:
        n["repoInfo"] = {"gitRepoDir": r.GitRepoDir, "gitRepoURL": r.GitRepoURL, "gitRepoBranch": r.GitRepoBranch, "targetPath": r.TargetPath}
    return n


def plan_to_node(plan) -> dict:
    """The plan as nested dicts in the reference's key order (no path conversion)."""
    inp, k8s = plan.Spec.Inputs, plan.Spec.Outputs.Kubernetes
    n = {}
    if plan.APIVersion:
        n["apiVersion"] = plan.APIVersion
    n["kind"] = plan.Kind
    if plan.Name:
        n["metadata"] = {"name": plan.Name}
    inputs = {"rootDir": inp.RootDir}
    if inp.K8sFiles:
        inputs["kubernetesYamls"] = list(inp.K8sFiles)
    if inp.QACaches:
        inputs["qaCaches"] = list(inp.QACaches)
    inputs["services"] = {name: [_service_node(s) for s in inp.Services[name]] for name in sorted(inp.Services)}
    if inp.TargetInfoArtifacts:
        inputs["targetInfoArtifacts"] = {k: list(inp.TargetInfoArtifacts[k]) for k in sorted(inp.TargetInfoArtifacts)}
    kub = {}
    if k8s.RegistryURL:
        kub["registryURL"] = k8s.RegistryURL
    if k8s.RegistryNamespace:
        kub["registryNamespace"] = k8s.RegistryNamespace
    kub["artifactType"] = k8s.ArtifactType
    tc = {}
    if k8s.TargetCluster.Type:
        tc["type"] = k8s.TargetCluster.Type
    if k8s.TargetCluster.Path:
        tc["path"] = k8s.TargetCluster.Path
    if tc:
        kub["targetCluster"] = tc
    if k8s.IgnoreUnsupportedKinds:
        kub["ignoreUnsupportedKinds"] = True
    n["spec"] = {"inputs": inputs, "outputs": {"kubernetes": kub}}
    return n


def dumps(plan, cwd=None) -> str:
    """WritePlan's bytes: a copy of the plan with absolute paths made relative (convertPathsEncode,
    planutils.go:146-163: m2kpath fields to rootDir / the assets' parent, rootDir itself to the
    working directory), encoded like common.WriteYaml."""
    p = _copy_plan(plan)
    _convert_paths(p, lambda x: GetRelativePath(p, x))
    p.Spec.Inputs.RootDir = _rel(cwd or os.getcwd(), p.Spec.Inputs.RootDir) if p.Spec.Inputs.RootDir else p.Spec.Inputs.RootDir
    out = []
    _emit(plan_to_node(p), 0, out)
    return "\n".join(out) + "\n"


def WritePlan(path: str, plan):
    """planutils.go:190-201.  Returns an error or None."""
    try:
        text = dumps(plan)
        with open(path, "w") as f:
            f.write(text)
    except (OSError, ValueError) as e:
        return e
    return None


# ---- decoding ----
def _bool(v) -> bool:
    return str(v).strip().lower() == "true"


def _strs(v):
    return [str(x) for x in (v or [])]


def _service_from_node(n) -> "plantypes.Service":
    r = n.get("repoInfo") or {}
    return plantypes.Service(
        ServiceName=str(n.get("serviceName", "")), ServiceRelPath=str(n.get("serviceRelPath", "")), Image=str(n.get("image", "")),
        TranslationType=str(n.get("translationType", "")), ContainerBuildType=str(n.get("containerBuildType", "")),
        SourceTypes=_strs(n.get("sourceType")), ContainerizationTargetOptions=_strs(n.get("targetOptions")),
        SourceArtifacts={str(k): _strs(v) for k, v in (n.get("sourceArtifacts") or {}).items()},
        BuildArtifacts={str(k): _strs(v) for k, v in (n.get("buildArtifacts") or {}).items()},
        UpdateContainerBuildPipeline=_bool(n.get("updateContainerBuildPipeline", False)),
        UpdateDeployPipeline=_bool(n.get("updateDeployPipeline", False)),
        RepoInfo=plantypes.RepoInfo(GitRepoDir=str(r.get("gitRepoDir", "")), GitRepoURL=str(r.get("gitRepoURL", "")),
                                    GitRepoBranch=str(r.get("gitRepoBranch", "")), TargetPath=str(r.get("targetPath", ""))))


def loads(text: str, where="<string>"):
    """ReadMove2KubeYaml + the decode into a Plan (utils.go:210-251), paths still as written.
    Scalars are read as strings (the reference decodes into typed fields, so `yes` or `1.0` in a string
    field stays a string); the two bool fields are converted explicitly."""
    node = yaml.load(text, Loader=yaml.BaseLoader)
    if not isinstance(node, dict):
        raise ValueError("%s: not a yaml mapping" % where)
    if "apiVersion" not in node:
        raise ValueError("Did not find apiVersion in the yaml file at path %s" % where)
    gv = str(node["apiVersion"])
    group = gv.rsplit("/", 1)[0] if "/" in gv else ""
    if group != GroupName:
        raise ValueError("The file at path %s doesn't have the correct group. Expected group %s Actual group %s" % (where, GroupName, group))
    p = plantypes.Plan(APIVersion=gv, Kind=str(node.get("kind", "")), Name=str((node.get("metadata") or {}).get("name", "")))
    spec = node.get("spec") or {}
    inp = spec.get("inputs") or {}
    p.Spec.Inputs = plantypes.Inputs(
        RootDir=str(inp.get("rootDir", "")), K8sFiles=_strs(inp.get("kubernetesYamls")), QACaches=_strs(inp.get("qaCaches")),
        Services={str(name): [_service_from_node(s) for s in (lst or [])] for name, lst in (inp.get("services") or {}).items()},
        TargetInfoArtifacts={str(k): _strs(v) for k, v in (inp.get("targetInfoArtifacts") or {}).items()})
    kub = (spec.get("outputs") or {}).get("kubernetes") or {}
    tc = kub.get("targetCluster") or {}
    p.Spec.Outputs = plantypes.Outputs(Kubernetes=plantypes.KubernetesOutput(
        RegistryURL=str(kub.get("registryURL", "")), RegistryNamespace=str(kub.get("registryNamespace", "")),
        ArtifactType=str(kub.get("artifactType", "")), TargetCluster=plantypes.TargetClusterType(Type=str(tc.get("type", "")), Path=str(tc.get("path", ""))),
        IgnoreUnsupportedKinds=_bool(kub.get("ignoreUnsupportedKinds", False))))
    return p


def ReadPlan(path: str, cwd=None):
    """planutils.go:164-176: decode, make rootDir absolute (against the working directory), then every
    m2kpath field absolute (convertPathsDecode, :130-144).  Returns (plan, err)."""
    try:
        with open(path) as f:
            p = loads(f.read(), path)
    except (OSError, ValueError, yaml.YAMLError) as e:
        return plantypes.Plan(), e
    p.Spec.Inputs.RootDir = os.path.normpath(os.path.join(cwd or os.getcwd(), p.Spec.Inputs.RootDir))
    _convert_paths(p, lambda x: GetAbsolutePath(p, x))
    return p, None


def diff_services(ours, theirs):
    """Differences between the services sections of two plans, both in wire form (relative paths):
    a list of (service name, index, field, ours, theirs).  Empty list = the detection results agree."""
    a = plan_to_node(ours)["spec"]["inputs"]["services"]
    b = plan_to_node(theirs)["spec"]["inputs"]["services"]
    out = []
    for name in sorted(set(a) | set(b)):
        la, lb = a.get(name, []), b.get(name, [])
        for i in range(max(len(la), len(lb))):
            sa, sb = (la[i] if i < len(la) else None), (lb[i] if i < len(lb) else None)
            if sa is None or sb is None:
                out.append((name, i, "<service>", "present" if sa else "missing", "present" if sb else "missing"))
                continue
            for k in sorted(set(sa) | set(sb)):
                if sa.get(k) != sb.get(k):
                    out.append((name, i, k, sa.get(k), sb.get(k)))
    return out
