"""The boundary data types of the path, mirrored from the reference's types/plan.

Reference: types/plan/plan.go:80-93,108-117 (enum strings), :186-215 (Service), :217-230 (NewService).
Only what Any2KubeTranslator.GetServiceOptions reads or writes is kept.
"""
from dataclasses import dataclass, field
from typing import Dict, List

# TranslationTypeValue / ContainerBuildTypeValue / SourceTypeValue strings (plan.go:80-117)
Any2KubeTranslation = "Any2Kube"
DockerFileContainerBuildTypeValue = "NewDockerfile"
ReuseDockerFileContainerBuildTypeValue = "ReuseDockerfile"
ReuseContainerBuildTypeValue = "Reuse"
CNBContainerBuildTypeValue = "CNB"
S2IContainerBuildTypeValue = "S2I"
ManualContainerBuildTypeValue = "Manual"
DirectorySourceTypeValue = "Directory"
SourceDirectoryArtifactType = "SourceCode"
SourceDirectoryBuildArtifactType = "SourceCode"


@dataclass
class RepoInfo:
    """plantypes.RepoInfo (plan.go:194-199): filled by Service.GatherGitInfo (go-git, host only)."""
    GitRepoDir: str = ""
    GitRepoURL: str = ""
    GitRepoBranch: str = ""
    TargetPath: str = ""


@dataclass
class Service:
    """plantypes.Service (plan.go:202-215)."""
    ServiceName: str = ""
    Image: str = ""
    TranslationType: str = ""
    ContainerBuildType: str = ""
    SourceTypes: List[str] = field(default_factory=list)
    ContainerizationTargetOptions: List[str] = field(default_factory=list)
    SourceArtifacts: Dict[str, List[str]] = field(default_factory=dict)
    BuildArtifacts: Dict[str, List[str]] = field(default_factory=dict)
    UpdateContainerBuildPipeline: bool = False
    UpdateDeployPipeline: bool = False
    ServiceRelPath: str = ""
    RepoInfo: "RepoInfo" = field(default_factory=RepoInfo)

    def AddSourceType(self, st: str) -> bool:
        if st not in self.SourceTypes:
            self.SourceTypes.append(st)
        return True


def service_to_plan_dict(s: "Service", root_dir: str, assets_path: str) -> dict:
    """One service the way WritePlan stores it (types/plan/planutils.go:164-201): the yaml keys of
    plan.go:202-215, with the `m2kpath` fields made relative (planutils.go:239-268) -- source and build
    artifacts to the plan's rootDir, Dockerfile/S2I target options to `m2kassets/...` when they live under
    the assets directory (builder names of CNB options are not paths and stay as they are)."""
    import os

    def rel_root(p):
        try:
            return os.path.relpath(p, root_dir)
        except ValueError:
            return p

    def rel_opt(p):
        if s.ContainerBuildType not in (DockerFileContainerBuildTypeValue, ReuseDockerFileContainerBuildTypeValue,
                                        S2IContainerBuildTypeValue):
            return p
        ap = os.path.abspath(assets_path)
        if os.path.abspath(p).startswith(ap + os.sep):
            return os.path.join("m2kassets", os.path.relpath(os.path.abspath(p), ap))
        return rel_root(p)

    return {
        "serviceName": s.ServiceName, "serviceRelPath": s.ServiceRelPath, "image": s.Image,
        "translationType": s.TranslationType, "containerBuildType": s.ContainerBuildType,
        "sourceType": list(s.SourceTypes), "targetOptions": [rel_opt(o) for o in s.ContainerizationTargetOptions],
        "sourceArtifacts": {k: [rel_root(x) for x in v] for k, v in s.SourceArtifacts.items()},
        "buildArtifacts": {k: [rel_root(x) for x in v] for k, v in s.BuildArtifacts.items()},
        "updateContainerBuildPipeline": s.UpdateContainerBuildPipeline, "updateDeployPipeline": s.UpdateDeployPipeline,
    }


def NewService(servicename: str, translationtype: str) -> Service:
    """plan.go:217-230."""
    return Service(ServiceName=servicename, Image=servicename + ":latest", TranslationType=translationtype,
                   SourceTypes=[], ContainerizationTargetOptions=[], SourceArtifacts={}, BuildArtifacts={},
                   UpdateDeployPipeline=False, ServiceRelPath="/" + servicename)


@dataclass
class Inputs:
    """plan.go:185-191"""
    RootDir: str = ""
    K8sFiles: List[str] = field(default_factory=list)
    QACaches: List[str] = field(default_factory=list)
    Services: Dict[str, List[Service]] = field(default_factory=dict)
    TargetInfoArtifacts: Dict[str, List[str]] = field(default_factory=dict)


@dataclass
class TargetClusterType:
    """plan.go:162-165"""
    Type: str = ""
    Path: str = ""


@dataclass
class KubernetesOutput:
    """plan.go:151-157"""
    RegistryURL: str = ""
    RegistryNamespace: str = ""
    ArtifactType: str = "Yamls"
    TargetCluster: "TargetClusterType" = field(default_factory=lambda: TargetClusterType(Type="Kubernetes"))
    IgnoreUnsupportedKinds: bool = False


@dataclass
class Outputs:
    Kubernetes: "KubernetesOutput" = field(default_factory=KubernetesOutput)


@dataclass
class Spec:
    Inputs: "Inputs" = field(default_factory=Inputs)
    Outputs: "Outputs" = field(default_factory=Outputs)


@dataclass
class Plan:
    """plantypes.Plan (plan.go:133-160); NewPlan's defaults (plan.go:400-426).  The wire format
    (ReadPlan / WritePlan, planutils.go:164-201) is in planio.py."""
    APIVersion: str = "move2kube.konveyor.io/v1alpha1"
    Kind: str = "Plan"
    Name: str = "myproject"
    Spec: "Spec" = field(default_factory=Spec)

    def SetRootDir(self, path: str):
        self.Spec.Inputs.RootDir = path

    def AddServicesToPlan(self, services: List[Service]):
        """plan.go:380-398 with merge rules :274-288 restricted to what this path produces:
        same name + same build type + same source dirs merge, anything else is appended."""
        for s in services:
            lst = self.Spec.Inputs.Services.setdefault(s.ServiceName, [])
            for old in lst:
                if (old.ContainerBuildType == s.ContainerBuildType and old.TranslationType == s.TranslationType
                        and old.SourceArtifacts.get(SourceDirectoryArtifactType) ==
                        s.SourceArtifacts.get(SourceDirectoryArtifactType)):
                    for o in s.ContainerizationTargetOptions:
                        if o not in old.ContainerizationTargetOptions:
                            old.ContainerizationTargetOptions.append(o)
                    break
            else:
                lst.append(s)


def NewPlan() -> Plan:
    return Plan()
