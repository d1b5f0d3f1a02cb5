"""CPU tests of the plan wire format (move2kube_b200/planio.py) against the reference's own golden plan
files (types/plan/planutils.go:164-201; goldens: internal/move2kube/testdata/expectedplanfornodejsapp.yaml
<- planner_test.go:99-128, internal/source/testdata/expectedservicesfor*.yaml <- any2kube_test.go:97-286),
stored as text in tests/golden/_reference_plans.json by tests/golden/make_golden.py."""
import json
import os
import subprocess
import sys

import pytest

from move2kube_b200 import containerizer, plan as plantypes, planio

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = json.load(open(os.path.join(REPO, "tests", "golden", "_reference_plans.json")))


@pytest.fixture()
def temp_assets(monkeypatch):
    monkeypatch.setattr(containerizer, "AssetsPath", "/tmp/m2k-1234/m2kassets")  # what CreateAssetsData sets (absolute)


def _nodejs_plan(root):
    p = plantypes.NewPlan()
    p.Name = "nodejs-app"
    p.Spec.Inputs.RootDir = root
    for bt, opts in (("NewDockerfile", ["/tmp/m2k-1234/m2kassets/dockerfiles/nodejs"]), ("S2I", ["/tmp/m2k-1234/m2kassets/s2i/nodejs"]),
                     ("CNB", ["cloudfoundry/cnb:cflinuxfs3", "gcr.io/buildpacks/builder"])):
        s = plantypes.NewService("nodejs", plantypes.Any2KubeTranslation)
        s.AddSourceType(plantypes.DirectorySourceTypeValue)
        s.ContainerBuildType = bt
        s.ContainerizationTargetOptions = list(opts)
        s.SourceArtifacts = {"SourceCode": [root]}
        s.BuildArtifacts = {"SourceCode": [root]}
        s.UpdateContainerBuildPipeline = s.UpdateDeployPipeline = True
        p.AddServicesToPlan([s])
    return p


def test_written_plan_is_the_reference_golden_byte_for_byte(temp_assets, tmp_path):
    """planner_test.go:99-128 compares CreatePlan(samples/nodejs) with this golden; written from the package
    directory internal/move2kube, rootDir is ../../samples/nodejs."""
    want = REF["planfornodejsapp"]["text"]
    p = _nodejs_plan("/src/samples/nodejs")
    assert planio.dumps(p, cwd="/src/internal/move2kube") == want
    out = tmp_path / "m2k.plan"
    assert planio.WritePlan(str(out), p) is None
    back, err = planio.ReadPlan(str(out))
    assert err is None and back == p


@pytest.mark.parametrize("key", sorted(REF))
def test_read_then_write_keeps_every_golden(temp_assets, tmp_path, key):
    text = REF[key]["text"]
    f = tmp_path / "golden.yaml"
    f.write_text(text)
    p, err = planio.ReadPlan(str(f), cwd="/work/pkg")
    assert err is None
    root = p.Spec.Inputs.RootDir
    assert os.path.isabs(root)
    for lst in p.Spec.Inputs.Services.values():
        for s in lst:
            for paths in list(s.SourceArtifacts.values()) + list(s.BuildArtifacts.values()):
                assert all(os.path.isabs(x) for x in paths)
            if s.ContainerBuildType in ("NewDockerfile", "S2I"):
                assert all(x.startswith("/tmp/m2k-1234/m2kassets/") for x in s.ContainerizationTargetOptions)
            else:
                assert s.ContainerizationTargetOptions == ["cloudfoundry/cnb:cflinuxfs3", "gcr.io/buildpacks/builder"]  # builder names are not paths
    again = planio.loads(planio.dumps(p, cwd="/work/pkg"))
    assert planio.plan_to_node(again) == planio.plan_to_node(planio.loads(text))
    assert planio.plan_to_node(planio.loads(text))["spec"]["inputs"]["services"] == _canon(REF[key]["services"])


def _canon(services):
    """The golden's services in planio's key order (three of the goldens were edited by hand: serviceRelPath last)."""
    order = ["serviceName", "serviceRelPath", "image", "translationType", "containerBuildType", "sourceType", "targetOptions",
             "sourceArtifacts", "buildArtifacts", "updateContainerBuildPipeline", "updateDeployPipeline", "repoInfo"]
    return {name: [{k: s[k] for k in order if k in s} for s in lst] for name, lst in sorted(services.items())}


def test_reader_checks_the_api_group_and_reads_typed_fields(tmp_path):
    with pytest.raises(ValueError, match="apiVersion"):
        planio.loads("kind: Plan\n")
    with pytest.raises(ValueError, match="correct group"):
        planio.loads("apiVersion: other.io/v1\nkind: Plan\n")
    _, err = planio.ReadPlan(str(tmp_path / "missing.plan"))
    assert err is not None
    p = planio.loads("apiVersion: move2kube.konveyor.io/v1alpha1\nkind: Plan\nspec:\n  inputs:\n    rootDir: /x\n    services:\n"
                     "      \"1.0\":\n        - serviceName: \"1.0\"\n          image: yes\n          updateDeployPipeline: true\n")
    s = p.Spec.Inputs.Services["1.0"][0]
    assert (s.ServiceName, s.Image, s.UpdateDeployPipeline, s.UpdateContainerBuildPipeline) == ("1.0", "yes", True, False)
    # and strings that would read back as something else are quoted on the way out
    text = planio.dumps(p, cwd="/")
    assert '"1.0":' in text and 'image: "yes"' in text
    assert planio.diff_services(planio.loads(text), p) == []


def test_paths_outside_the_assets_directory_follow_the_root(temp_assets):
    p = _nodejs_plan("/a/b/root")
    svc = p.Spec.Inputs.Services["nodejs"][0]
    svc.ContainerizationTargetOptions = ["/a/b/root/custom/dockerfiles/mine"]   # a user-supplied option directory
    svc.SourceArtifacts["Other"] = ["/not/converted"]                           # not one of the m2kpath keys
    n = planio.plan_to_node(planio.loads(planio.dumps(p, cwd="/a")))
    got = n["spec"]["inputs"]["services"]["nodejs"][0]
    assert n["spec"]["inputs"]["rootDir"] == "b/root"
    assert got["targetOptions"] == ["custom/dockerfiles/mine"]
    assert got["sourceArtifacts"] == {"Other": ["/not/converted"], "SourceCode": ["."]}


def test_plan_diff_tool(temp_assets, tmp_path):
    a, b = tmp_path / "a.plan", tmp_path / "b.plan"
    a.write_text(REF["planfornodejsapp"]["text"])
    p = _nodejs_plan("/src/samples/nodejs")
    p.Spec.Inputs.Services["nodejs"] = p.Spec.Inputs.Services["nodejs"][:2]       # what a box without Docker produces
    b.write_text(planio.dumps(p, cwd="/elsewhere"))
    tool = os.path.join(REPO, "tools", "plan_diff.py")
    r = subprocess.run([sys.executable, tool, str(b), str(a)], stdout=subprocess.PIPE, text=True)
    assert r.returncode == 1 and "nodejs[2].<service>" in r.stdout
    r = subprocess.run([sys.executable, tool, str(b), str(a), "--ignore-cnb"], stdout=subprocess.PIPE, text=True)
    assert r.returncode == 0 and "0 difference(s)" in r.stdout
