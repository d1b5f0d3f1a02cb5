"""GPU tests of the host-side mirror of the reference's interfaces (pytest -m gpu): they read like
internal/source/any2kube_test.go and internal/common/utils_test.go, with the golden fixtures as
expected values."""
import os

import pytest

import golden_check as gc
import treespec
from move2kube_b200 import common, containerizer, detection, plan as plantypes, source

pytestmark = pytest.mark.gpu

GOLDENS = {g["name"]: g for g in gc.load_goldens()}
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture()
def assets(tmp_path, monkeypatch):
    """An assets directory with the 18 built-in option directories (the mirror only looks at where
    the detect scripts are, it never runs the built-in ones)."""
    root = tmp_path / "m2kassets"
    for n in ["django", "golang", "java-war-jboss", "java-war-liberty", "java-war-tomcat", "javaant", "javagradle",
              "javamaven", "nodejs", "php", "python", "ruby"]:
        (root / "dockerfiles" / n).mkdir(parents=True)
        (root / "dockerfiles" / n / "m2kdfdetect.sh").write_text("#!/bin/sh\nexit 1\n")
    for n in ["golang", "java", "nodejs", "php", "python", "ruby"]:
        (root / "s2i" / n).mkdir(parents=True)
        (root / "s2i" / n / "m2ks2idetect.sh").write_text("#!/bin/sh\nexit 1\n")
    monkeypatch.setattr(containerizer, "AssetsPath", str(root))
    yield str(root)
    detection.reset()


def golden_services(spec, root, assets, flavor="bash"):
    out = []
    for rel, name, bt, topts in spec["expect"][flavor]["services"]:
        path = root if rel == "." else os.path.join(root, rel)
        opts = []
        for o in topts:
            if o.startswith("m2kassets/"):
                opts.append(os.path.join(assets, o[len("m2kassets/"):]))
            else:
                opts.append(os.path.join(root, o[len("tree:"):]))
        out.append((name, bt, path, opts))
    return out


@pytest.mark.parametrize("name", ["samples", "samples_nodejs", "ignorecase1", "ignorecase2", "ignorecontents", "multiignores",
                                  "emptydir", "nasty", "userscript", "random007", "random031"])
def test_get_service_options(tmp_path, assets, name):
    spec = GOLDENS[name]
    root = treespec.materialize(spec, str(tmp_path / name))
    translator = source.Any2KubeTranslator()
    plan = plantypes.NewPlan()
    plan.SetRootDir(root)
    services, err = translator.GetServiceOptions(root, plan)
    assert err is None
    got = [(s.ServiceName, s.ContainerBuildType, s.SourceArtifacts["SourceCode"][0], s.ContainerizationTargetOptions)
           for s in services]
    assert got == golden_services(spec, root, assets)
    for s in services:
        assert s.Image == s.ServiceName + ":latest" and s.ServiceRelPath == "/" + s.ServiceName
        assert s.TranslationType == "Any2Kube" and s.SourceTypes == ["Directory"]
        assert s.BuildArtifacts["SourceCode"] == s.SourceArtifacts["SourceCode"]
        assert s.UpdateContainerBuildPipeline and s.UpdateDeployPipeline


def test_nonexistent_directory_gives_no_services_and_no_error(tmp_path, assets):
    """any2kube_test.go:38-53"""
    services, err = source.Any2KubeTranslator().GetServiceOptions(str(tmp_path / "nope"), plantypes.NewPlan())
    assert services == [] and err is None


def test_source_directory_already_in_the_plan(tmp_path, assets):
    """any2kube_test.go:179-212: a directory some other translator already containerized is skipped."""
    spec = GOLDENS["samples_nodejs_pre"]
    root = treespec.materialize(spec, str(tmp_path / "nodejs"))
    plan = plantypes.NewPlan()
    svc = plantypes.NewService("nodejs", "Dockerfile2Kube")
    svc.SourceArtifacts["SourceCode"] = [root]
    plan.AddServicesToPlan([svc])
    services, err = source.Any2KubeTranslator().GetServiceOptions(root, plan)
    assert services == [] and err is None


def test_without_assets_only_host_options_remain(tmp_path, monkeypatch):
    """The reference's own unit tests run with a non-existent AssetsPath: no
    Dockerfile / S2I options exist then, and without a container runtime no CNB ones either."""
    monkeypatch.setattr(containerizer, "AssetsPath", str(tmp_path / "no-assets"))
    spec = GOLDENS["samples_nodejs"]
    root = treespec.materialize(spec, str(tmp_path / "nodejs"))
    services, err = source.Any2KubeTranslator().GetServiceOptions(root, plantypes.NewPlan())
    assert services == [] and err is None
    # with a CNB provider hooked in, the golden of any2kube_test.go:97-131 comes back
    monkeypatch.setattr(containerizer.CNBContainerizer, "provider", staticmethod(lambda path, builder: True))
    containerizer.CNBContainerizer._cache.clear()
    services, err = source.Any2KubeTranslator().GetServiceOptions(root, plantypes.NewPlan())
    assert [(s.ServiceName, s.ContainerBuildType, s.ContainerizationTargetOptions) for s in services] == \
        [("nodejs", "CNB", ["cloudfoundry/cnb:cflinuxfs3", "gcr.io/buildpacks/builder"])]
    containerizer.CNBContainerizer._cache.clear()
    detection.reset()


def test_get_files_by_name_and_ext(tmp_path, assets):
    """utils_test.go:32-140: Walk order, error on a missing root."""
    spec = GOLDENS["multiignores"]
    root = treespec.materialize(spec, str(tmp_path / "mi"))
    files, err = common.GetFilesByName(root, [".m2kignore"])
    assert err is None
    assert files == [os.path.join(root, p) for p in [".m2kignore", "a/a/.m2kignore", "b/.m2kignore"]]
    _, err = common.GetFilesByName(str(tmp_path / "nope"), [".m2kignore"])
    assert err is not None
    spec = GOLDENS["samples"]
    root = treespec.materialize(spec, str(tmp_path / "samples"))
    files, err = common.GetFilesByExt(root, [".yml", ".yaml"])
    want = []
    for dirpath, dirnames, filenames in os.walk(root):
        for fn in filenames:
            if os.path.splitext(fn)[1] in (".yml", ".yaml"):
                want.append(os.path.join(dirpath, fn))
    assert err is None and sorted(files) == sorted(want) and files == sorted(files, key=lambda p: p.split(os.sep))
    files, err = common.GetFilesByName(root, ["package.json", "pom.xml"])
    want = []
    for dirpath, dirnames, filenames in os.walk(root):
        want += [os.path.join(dirpath, fn) for fn in filenames if fn in ("package.json", "pom.xml")]
    want.sort(key=lambda p: p.split(os.sep))
    assert err is None and files == want and len(files) >= 3
    assert common.IsStringPresent(["/A/b", "/c"], "/a/B") and not common.IsStringPresent(["/a"], "/ab")


def test_dockerfile_translator_finds_dockerfiles(tmp_path, assets):
    d = tmp_path / "df"
    (d / "svc").mkdir(parents=True)
    (d / "svc" / "Dockerfile").write_text("FROM alpine\nRUN true\n")
    (d / "svc" / "notes.txt").write_text("FROM is just the first word here, twice\n")
    (d / "build.Dockerfile").write_text("ARG V=1\nFROM alpine:$V AS b\n")
    (d / "broken").write_text("FROM alpine\nENV X\n")
    # filepath.Walk interleaves files and sub-directories by name: a file sorted after "svc" comes after svc's files
    (d / "zz.Dockerfile").write_text("FROM scratch\n")
    files, err = source.DockerfileTranslator().GetDockerfiles(str(d))
    assert err is None
    assert [os.path.relpath(f, str(d)) for f in files] == ["build.Dockerfile", "svc/Dockerfile", "zz.Dockerfile"]


def test_translate_time_detect_is_answered_from_the_pass(tmp_path, assets):
    """GetContainer re-runs the option's detect script to read its JSON (dockerfilecontainerizer.go:105-129,
    s2icontainerizer.go:97-115): the mirror's detect() answers from the detection pass instead, with the
    bytes the reference's script prints (internal/assets/dockerfiles/python/m2kdfdetect.sh:21-24)."""
    import json
    app = tmp_path / "svc"
    (app / "pkg").mkdir(parents=True)
    (app / "requirements.txt").write_text("flask\n")
    (app / "pkg" / "run.py").write_text("import x\nif __name__ == '__main__':\n    x.go()\n")
    (app / "lib.war").write_text("not really a war")
    cs = containerizer.Containerizers()
    cs.InitContainerizers(str(app))
    df, s2i = cs.containerizers[0], cs.containerizers[1]
    out, err = df.detect(os.path.join(assets, "dockerfiles", "python"), str(app))
    assert err is None and out == '{"main_script_rel_path": "pkg/run.py", "app_name": "app", "port": 8080}'
    out, err = s2i.detect(os.path.join(assets, "s2i", "python"), str(app))
    assert err is None and json.loads(out)["app_file"] == "pkg/run.py" and json.loads(out)["builder"].endswith("python-36-rhel7:latest")
    out, err = df.detect(os.path.join(assets, "dockerfiles", "java-war-liberty"), str(app))
    assert err is None and out == '{"port":9080, "war_path":"lib.war"}'
    out, err = df.detect(os.path.join(assets, "dockerfiles", "nodejs"), str(app))
    assert err is not None and out == ""          # the script would print nothing and exit 1
    cs.close()


def test_a_root_that_is_not_a_directory(tmp_path, assets):
    """filepath.Walk Lstats its root: a symlink to a directory, or a file, is visited alone.  Any2Kube
    returns no services (any2kube.go:63-65); getDockerfileServices asks isDockerFile about the root
    itself (dockerfile2kube.go:154-167)."""
    real = tmp_path / "real"
    real.mkdir()
    (real / "package.json").write_text("{}")
    (real / "Dockerfile").write_text("FROM node:14\nCMD [\"node\"]\n")
    link = tmp_path / "link"
    os.symlink(str(real), str(link))
    services, err = source.Any2KubeTranslator().GetServiceOptions(str(real), plantypes.NewPlan())
    assert err is None and len(services) >= 1
    services, err = source.Any2KubeTranslator().GetServiceOptions(str(link), plantypes.NewPlan())
    assert err is None and services == []
    services, err = source.Any2KubeTranslator().GetServiceOptions(str(real / "package.json"), plantypes.NewPlan())
    assert err is None and services == []
    tr = source.DockerfileTranslator()
    assert tr.GetDockerfiles(str(real)) == ([str(real / "Dockerfile")], None)
    assert tr.GetDockerfiles(str(link)) == ([], None)                      # reading a directory fails, so does Parse
    assert tr.GetDockerfiles(str(real / "Dockerfile")) == ([str(real / "Dockerfile")], None)
    assert tr.GetDockerfiles(str(real / "package.json")) == ([], None)
    flink = tmp_path / "df-link"
    os.symlink(str(real / "Dockerfile"), str(flink))
    assert tr.GetDockerfiles(str(flink)) == ([str(flink)], None)            # os.Open follows the link
    files, err = tr.GetDockerfiles(str(tmp_path / "nope"))
    assert files == [] and err is not None


@pytest.mark.parametrize("key", ["planfornodejsapp", "servicesfornodejsapp", "servicesfornodejsappwithm2kignorecase1",
                                 "servicesforjavamavenappwithm2kignorecase2"])
def test_against_the_reference_golden_plans(tmp_path, tmp_path_factory, assets, monkeypatch, key):
    """The reference's OWN golden YAML files for this path (planner_test.go:99-128, any2kube_test.go:97-286),
    field by field in the plan's wire format.  They were recorded with Docker available, so the CNB
    provider hook stands in for the builders' detector; the three any2kube goldens were recorded without the
    assets directory (only the CNB option), the planner golden with it."""
    ref = gc.reference_plans()[key]
    spec = GOLDENS[ref["tree"]]
    # the service is named after the directory: samples/nodejs in the reference
    # (the sources live outside the directory that holds the assets: common.TempPath never contains the input tree)
    src_base = tmp_path_factory.mktemp("sources")
    root = treespec.materialize(spec, str(src_base / ("nodejs" if ref["tree"] == "samples_nodejs" else ref["tree"])))
    if key != "planfornodejsapp":
        monkeypatch.setattr(containerizer, "AssetsPath", str(tmp_path / "no-assets"))
    # stand-in for the builder images' detector: what the goldens pin is that both builders accept a
    # directory holding package.json / pom.xml and nothing else in these trees
    monkeypatch.setattr(containerizer.CNBContainerizer, "provider", staticmethod(
        lambda path, builder: any(os.path.isfile(os.path.join(path, m)) for m in ("package.json", "pom.xml"))))
    containerizer.CNBContainerizer._cache.clear()
    services, err = source.Any2KubeTranslator().GetServiceOptions(root, plantypes.NewPlan())
    assert err is None
    got = {}
    for s in services:
        got.setdefault(s.ServiceName, []).append(plantypes.service_to_plan_dict(s, root, containerizer.AssetsPath))
    assert got == ref["services"]
    if key == "planfornodejsapp":
        # and as a FILE: the plan these services make, written by planio.WritePlan's encoder from the package
        # directory the reference's test runs in, is the reference's golden file byte for byte
        from move2kube_b200 import planio
        p = plantypes.NewPlan()
        p.Name = "nodejs-app"
        p.Spec.Inputs.RootDir = root
        p.AddServicesToPlan(services)
        cwd = os.path.join(os.path.dirname(os.path.dirname(root)), "internal", "move2kube")  # rootDir: ../../samples/nodejs
        text = planio.dumps(p, cwd=cwd)
        want = ref["text"].replace("rootDir: ../../samples/nodejs", "rootDir: " + os.path.relpath(root, cwd))
        assert text == want
    containerizer.CNBContainerizer._cache.clear()


def test_a_result_may_outlive_its_context(tmp_path):
    """A host with a garbage collector releases in any order: a result freed after m2k_ctx_destroy gives nothing
    back to the (gone) context's pools and frees its buffers itself."""
    from move2kube_b200 import _abi, lib
    c = lib.Context(0)
    p = _abi.SynthParams(seed=3, n_files=300, n_dirs=30, max_depth=4, n_ignore_files=2, total_bytes=4 << 20)
    tree = c.synth_tree(p, host_bytes=True)
    res = c.detect(tree)
    n = res.contents.n_services
    L = c.L
    c.close()                      # the context goes first
    assert res.contents.n_services == n and res.contents.dir_opts[0] is not None   # the result's arrays are still there
    L.m2k_result_free(res)         # ... and can be released now
    L.m2k_tree_free(None, tree)
    c2 = lib.Context(0)            # a new context works as usual
    t2 = c2.synth_tree(p, host_bytes=True)
    r2 = c2.detect(t2)
    assert r2.contents.n_services == n
    c2.result_free(r2)
    c2.tree_free(t2)
    c2.close()


@pytest.mark.parametrize("name", ["samples", "pairwise", "nasty"])
def test_the_c_example_prints_what_the_binding_sees(tmp_path, name):
    """examples/detect.c drives the C ABI from plain C (m2k_pack_detect, m2k_option_json): its lines must be the ones
    the ctypes binding gets from the same calls -- which the other tests hold against the goldens."""
    import subprocess
    from test_host import _build_c_example
    from move2kube_b200 import _abi, lib
    root = str(tmp_path / "tree")
    treespec.materialize(GOLDENS[name], root)
    exe = _build_c_example(tmp_path)
    p = subprocess.run([exe, root], capture_output=True)
    assert p.returncode == 0, p.stderr
    got = p.stdout.split(b"\n")[:-1]
    ctx = lib.Context(0)
    tp, rp = ctx.pack_detect(os.fsencode(root))
    t, r = tp.contents, rp.contents
    names = bytes(t.names[0:t.entry_name_off[t.n_entries]])

    def rel(d):
        parts = []
        while t.dir_parent[d] != _abi.NONE:
            e = t.dir_entry[d]
            parts.append(names[t.entry_name_off[e]:t.entry_name_off[e + 1]])
            d = t.dir_parent[d]
        return b"/".join(reversed(parts)) or b"."
    want = []
    for k in range(r.n_services):
        d = r.service_dirs[k]
        for opt in range(18):
            if r.dir_opts[d] >> opt & 1:
                want.append(rel(d) + b"\t%d\t" % opt + lib.option_json(tp, rp, d, opt).rstrip(b"\n"))
    for f in range(r.n_files):
        if r.file_flags[f] & _abi.FILE_DOCKERFILE:
            e = t.file_entry[f]
            want.append(b"dockerfile\t" + rel(t.entry_dir[e]) + b"/" + names[t.entry_name_off[e]:t.entry_name_off[e + 1]])
    assert got == want
    assert len(want) >= (3 if name == "samples" else 20)
    ctx.result_free(rp)
    ctx.tree_free(tp)
    ctx.close()
