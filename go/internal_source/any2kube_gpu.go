// Drop-in for internal/source/any2kube.go of the reference (package source: the Translator interface
// has the unexported method newService, so an implementation must live in that package --
// internal/source/translator.go:27-32).  NOT COMPILED HERE (no Go toolchain on this image); it shows
// the wiring a maintainer adds.  Registration: replace `new(Any2KubeTranslator)` by
// `new(Any2KubeGPUTranslator)` in GetSourceLoaders (translator.go:35-39), e.g. behind an opt-in
// environment variable M2K_GPU_DETECT=1 so defaults do not change.
package source

import (
	"os"
	"path/filepath"
	"strconv"
	"strings"

	"github.com/konveyor/move2kube/internal/common"
	"github.com/konveyor/move2kube/internal/containerizer"
	"github.com/konveyor/move2kube/internal/m2kdetect" // go/m2kdetect of this repository
	irtypes "github.com/konveyor/move2kube/internal/types"
	plantypes "github.com/konveyor/move2kube/types/plan"
	log "github.com/sirupsen/logrus"
)

// Any2KubeGPUTranslator implements Translator with the detection pass on the GPU.
type Any2KubeGPUTranslator struct {
	Any2KubeTranslator // Translate() and newService() are inherited unchanged (any2kube.go:108-149)
}

// GetServiceOptions mirrors any2kube.go:43-105.
func (t *Any2KubeGPUTranslator) GetServiceOptions(inputPath string, plan plantypes.Plan) ([]plantypes.Service, error) {
	services := []plantypes.Service{}
	preContainerizedSourcePaths := []string{}
	for _, existingServices := range plan.Spec.Inputs.Services {
		for _, existingService := range existingServices {
			if len(existingService.SourceArtifacts[plantypes.SourceDirectoryArtifactType]) > 0 {
				preContainerizedSourcePaths = append(preContainerizedSourcePaths, existingService.SourceArtifacts[plantypes.SourceDirectoryArtifactType][0])
			}
		}
	}
	if info, err := os.Lstat(inputPath); err != nil {
		// filepath.Walk would hand this error to the callback, which warns and returns nil (any2kube_test.go:38-53)
		log.Warnf("Skipping path %q due to error. Error: %q", inputPath, err)
		return services, nil
	} else if !info.IsDir() {
		// filepath.Walk Lstats its root: a file, or a symlink even to a directory, is visited alone and
		// `!info.IsDir()` returns nil (any2kube.go:63-65)
		return services, nil
	}
	// M2K_DEVICES="0,1,2,3": use several GPUs of the box from this one process (m2k_detect_multi); default: device 0
	var res *m2kdetect.Result
	var err error
	if devs := parseDevices(os.Getenv("M2K_DEVICES")); len(devs) > 1 {
		res, err = m2kdetect.DetectMulti(inputPath, preContainerizedSourcePaths, os.Getenv("M2K_SH") == "dash", devs)
	} else {
		res, err = m2kdetect.Detect(inputPath, preContainerizedSourcePaths, os.Getenv("M2K_SH") == "dash")
	}
	if err != nil {
		// no CPU fallback: a GPU failure is an error of this translator, never silently empty results
		log.Errorf("GPU source detection failed for %q Error: %q", inputPath, err)
		return services, err
	}
	containerizers := new(containerizer.Containerizers)
	containerizers.InitContainerizersGPU(inputPath, res) // containerizers_gpu.go
	emit := func(d uint32) {
		path := res.DirPath[d]
		for _, containerizationOption := range containerizers.GetContainerizationOptions(plan, path) {
			service := t.newService(filepath.Base(path))
			service.ContainerBuildType = containerizationOption.ContainerizationType
			service.ContainerizationTargetOptions = containerizationOption.TargetOptions
			if !common.IsStringPresent(service.BuildArtifacts[plantypes.SourceDirectoryBuildArtifactType], path) {
				service.SourceArtifacts[plantypes.SourceDirectoryArtifactType] = append(service.SourceArtifacts[plantypes.SourceDirectoryArtifactType], path)
				service.BuildArtifacts[plantypes.SourceDirectoryBuildArtifactType] = append(service.BuildArtifacts[plantypes.SourceDirectoryBuildArtifactType], path)
			}
			if foundRepo, err := service.GatherGitInfo(path, plan); foundRepo && err != nil {
				log.Warnf("Error while parsing the git repo at path %q Error: %q", path, err)
			}
			services = append(services, service)
		}
	}
	if !containerizers.NeedsHostOptions() {
		// the walk ran on the device (k_walk_local + k_walk_reach): directories in state SERVICE, in walk order
		for _, d := range res.ServiceDirs {
			emit(d)
		}
		return services, nil
	}
	// user-supplied detect scripts or CNB: replay the walk over the packed tree, asking the host
	// containerizers per visited directory; ignore flags and built-in option bits stay the device's
	for d := uint32(0); int(d) < len(res.DirPath); {
		ign := res.DirIgnore[d]
		prune := false
		switch {
		case res.DirState[d] == m2kdetect.DirWalkError:
			log.Warnf("Skipping path %q due to error.", res.DirPath[d])
			prune = true
		case ign&4 != 0:
			prune = true
		case ign&1 != 0:
			prune = ign&2 != 0
		default:
			if len(containerizers.GetContainerizationOptions(plan, res.DirPath[d])) == 0 {
				prune = ign&2 != 0
			} else {
				emit(d)
				prune = true
			}
		}
		if prune {
			d = res.DirEnd[d]
		} else {
			d++
		}
	}
	return services, nil
}

var _ Translator = (*Any2KubeGPUTranslator)(nil)
var _ = irtypes.IR{}

// parseDevices reads a comma-separated list of CUDA device numbers.
func parseDevices(v string) []int {
	devs := []int{}
	for _, f := range strings.Split(v, ",") {
		if n, err := strconv.Atoi(strings.TrimSpace(f)); err == nil && n >= 0 {
			devs = append(devs, n)
		}
	}
	return devs
}
