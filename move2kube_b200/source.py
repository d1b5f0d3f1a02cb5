"""Mirror of the source translators' detection half (reference: internal/source/translator.go:27-39,
internal/source/any2kube.go:35-105,143-181, internal/source/dockerfile2kube.go:117-170).

`Any2KubeTranslator.GetServiceOptions(inputPath, plan)` keeps the reference's contract: a missing
root yields ([], None) (any2kube_test.go:38-53), walk problems are logged and skipped, the only
error returned is the walk's own -- plus, here, a GPU failure (there is no CPU fallback, so it must
surface: SURVEY.md section 8b).
"""
import logging
import os
import stat

from . import _abi, containerizer, detection, lib, plan as plantypes

log = logging.getLogger("move2kube_b200")


class Any2KubeTranslator:
    """Implements Translator for any source folder that can be containerized (any2kube.go:35-41)."""

    def GetTranslatorType(self):
        return plantypes.Any2KubeTranslation

    def newService(self, serviceName: str) -> plantypes.Service:
        """any2kube.go:143-149"""
        service = plantypes.NewService(serviceName, self.GetTranslatorType())
        service.AddSourceType(plantypes.DirectorySourceTypeValue)
        service.UpdateContainerBuildPipeline = True
        service.UpdateDeployPipeline = True
        return service

    def getIgnorePaths(self, inputPath: str, det: "detection.Detection" = None):
        """any2kube.go:151-181 -- answered from the detection pass (the .m2kignore files were found by
        the name-feature kernel and parsed once)."""
        own = det is None
        det = det or detection.detect(inputPath, skip_content=True)
        try:
            return ([os.fsdecode(p) for p in det.ignore_dirs], [os.fsdecode(p) for p in det.ignore_contents])
        finally:
            if own:
                det.close()

    def GetServiceOptions(self, inputPath: str, plan: plantypes.Plan):
        """any2kube.go:43-105.  Returns (services, err)."""
        services = []
        preContainerizedSourcePaths = []
        for existingServices in plan.Spec.Inputs.Services.values():
            for existingService in existingServices:
                src = existingService.SourceArtifacts.get(plantypes.SourceDirectoryArtifactType, [])
                if len(src) > 0:
                    preContainerizedSourcePaths.append(src[0])
        if not os.path.lexists(inputPath):
            # filepath.Walk hands the Lstat error to the callback, which warns and returns nil
            log.warning("Skipping path %r due to error. Error: no such file or directory", inputPath)
            return services, None
        if not stat.S_ISDIR(os.lstat(inputPath).st_mode):
            # filepath.Walk Lstats the root: a file, or a symlink (even to a directory), is handed to the
            # callback once and `!info.IsDir()` returns nil (any2kube.go:63-65) -- no services
            return services, None
        try:
            det = detection.detect(inputPath, pre=preContainerizedSourcePaths, cache=False)
        except lib.M2KError as e:
            log.error("GPU source detection failed for %r: %s", inputPath, e)
            return services, e
        try:
            containerizers = containerizer.Containerizers()
            containerizers.InitContainerizers(inputPath, det)
            builtin_only = not containerizers.has_user_options() and self._assets_complete(containerizers)
            if builtin_only:
                # the walk's decisions were made on the device (k_walk_local + k_walk_reach): emit in walk order
                for d in det.service_dirs:
                    self._emit(services, containerizers, plan, det, d)
            else:
                self._host_walk(services, containerizers, plan, det)
        finally:
            det.close()
        return services, None

    # -- helpers
    @staticmethod
    def _assets_complete(containerizers) -> bool:
        """True when every built-in option directory was found under AssetsPath, i.e. the device's
        dir_opts bits and the registered options describe the same set."""
        n = 0
        for c in containerizers.containerizers:
            if isinstance(c, containerizer._ScriptContainerizer):
                n += sum(1 for o in c.options if containerizer._builtin_index(o, c.kind) is not None)
        return n == _abi.N_DF + _abi.N_S2I

    def _emit(self, services, containerizers, plan, det, d: int):
        path = os.fsdecode(det.dir_path[d])
        for containerizationOption in containerizers.GetContainerizationOptions(plan, path):
            serviceName = os.path.basename(path.rstrip("/")) or path
            service = self.newService(serviceName)
            service.ContainerBuildType = containerizationOption.ContainerizationType
            service.ContainerizationTargetOptions = containerizationOption.TargetOptions
            service.SourceArtifacts.setdefault(plantypes.SourceDirectoryArtifactType, []).append(path)
            service.BuildArtifacts.setdefault(plantypes.SourceDirectoryBuildArtifactType, []).append(path)
            services.append(service)

    def _host_walk(self, services, containerizers, plan, det):
        """The walk of any2kube.go:58-98 replayed over the packed tree when some options need the host
        (user-supplied detect scripts, a CNB provider, or an incomplete assets directory): the
        per-directory ignore flags and built-in option bits still come from the device."""
        d = 0
        n = det.n_dirs
        while d < n:
            ign = det.dir_ignore[d]
            prune = False
            if det.dir_flags[d] & 1:
                log.warning("Skipping path %r due to error.", det.dir_path[d])
                prune = True
            elif ign & 4:
                prune = True
            elif ign & 1:
                prune = bool(ign & 2)
            else:
                path = os.fsdecode(det.dir_path[d])
                cops = containerizers.GetContainerizationOptions(plan, path)
                if len(cops) == 0:
                    prune = bool(ign & 2)
                else:
                    self._emit(services, containerizers, plan, det, d)
                    prune = True
            d = det.dir_end[d] if prune else d + 1


class DockerfileTranslator:
    """The detection half of Dockerfile2Kube (dockerfile2kube.go:117-170): which files of the tree are
    Dockerfiles.  Bucketing them into services (:171-263) is host string logic downstream of the path."""

    def GetDockerfiles(self, inputPath: str):
        """getDockerfileServices' walk (dockerfile2kube.go:146-170).  Returns (paths, err)."""
        try:
            st = os.stat(inputPath)
        except FileNotFoundError as e:  # os.Stat + IsNotExist (:147-149)
            return [], e
        except OSError:
            st = None
        if st is None or not stat.S_ISDIR(os.lstat(inputPath).st_mode):
            # the Walk visits the root alone, as a non-directory: isDockerFile(root).  os.Open follows a
            # symlink; reading a directory (a symlink to one) fails, and so does Parse
            try:
                with open(inputPath, "rb") as f:
                    data = f.read()
            except OSError:
                return [], None
            try:
                return ([inputPath] if detection.context().is_dockerfile(data) else []), None
            except lib.M2KError as e:
                return [], e
        det = detection.detect(inputPath, cache=False)
        try:
            return [os.fsdecode(p) for p in det.dockerfiles()], None
        finally:
            det.close()


def GetSourceLoaders():
    """translator.go:35-39 lists all translators; the two whose detection runs on the GPU are here."""
    return [DockerfileTranslator(), Any2KubeTranslator()]
