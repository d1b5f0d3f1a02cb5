// Drop-in additions to internal/containerizer of the reference (package containerizer).  NOT COMPILED
// HERE (no Go toolchain on this image).  The Containerizer interface (containerizer.go:37-42) is kept:
// Init / GetTargetOptions / GetContainer / GetContainerBuildStrategy.  Only GetTargetOptions changes --
// for option directories that are the built-in assets it reads the option bit the GPU computed for
// the directory instead of fork/exec'ing the detect script; everything else (GetContainer, i.e.
// translate-time rendering, user-supplied option directories, CNB, Reuse) is the reference's code.
package containerizer

import (
	"path/filepath"
	"sort"
	"strings"

	"github.com/konveyor/move2kube/internal/common"
	"github.com/konveyor/move2kube/internal/m2kdetect"
	plantypes "github.com/konveyor/move2kube/types/plan"
)

type gpuScriptContainerizer struct {
	res      *m2kdetect.Result
	dirIndex map[string]uint32
	names    []string // built-in option names in bit order
	kind     string   // "dockerfiles" | "s2i"
	shift    uint     // bit position of option 0 in DirOpts
}

func (g *gpuScriptContainerizer) builtinIndex(optionDir string) int {
	if filepath.Base(filepath.Dir(optionDir)) != g.kind {
		return -1
	}
	for i, n := range g.names {
		if n == filepath.Base(optionDir) {
			return i
		}
	}
	return -1
}

// DockerfileGPUContainerizer wraps the reference's DockerfileContainerizer.
type DockerfileGPUContainerizer struct {
	DockerfileContainerizer
	gpu gpuScriptContainerizer
}

// GetTargetOptions mirrors dockerfilecontainerizer.go:62-74.
func (d *DockerfileGPUContainerizer) GetTargetOptions(plan plantypes.Plan, path string) []string {
	targetOptions := []string{}
	id, known := d.gpu.dirIndex[path]
	for _, dfcontainerizer := range d.dfcontainerizers {
		if i := d.gpu.builtinIndex(dfcontainerizer); i >= 0 && known {
			if d.gpu.res.DirOpts[id]>>(d.gpu.shift+uint(i))&1 == 1 {
				targetOptions = append(targetOptions, dfcontainerizer)
			}
			continue
		}
		if _, err := d.detect(dfcontainerizer, path); err == nil { // user-supplied option: host exec, as the reference
			targetOptions = append(targetOptions, dfcontainerizer)
		}
	}
	return targetOptions
}

// S2IGPUContainerizer wraps the reference's S2IContainerizer (s2icontainerizer.go:63-75).
type S2IGPUContainerizer struct {
	S2IContainerizer
	gpu gpuScriptContainerizer
}

// GetTargetOptions mirrors s2icontainerizer.go:63-75.
func (s *S2IGPUContainerizer) GetTargetOptions(plan plantypes.Plan, path string) []string {
	targetOptions := []string{}
	id, known := s.gpu.dirIndex[path]
	for _, s2icontainerizer := range s.s2icontainerizers {
		if i := s.gpu.builtinIndex(s2icontainerizer); i >= 0 && known {
			if s.gpu.res.DirOpts[id]>>(s.gpu.shift+uint(i))&1 == 1 {
				targetOptions = append(targetOptions, s2icontainerizer)
			}
			continue
		}
		if _, err := s.detect(s2icontainerizer, path); err == nil {
			targetOptions = append(targetOptions, s2icontainerizer)
		}
	}
	return targetOptions
}

// InitContainerizersGPU mirrors InitContainerizers (containerizer.go:56-62): same four containerizers,
// same order, Init(path) then Init(common.AssetsPath); the user tree's option directories come from
// the name-feature bits of the detection pass instead of two more GetFilesByName walks.
func (c *Containerizers) InitContainerizersGPU(path string, res *m2kdetect.Result) {
	idx := make(map[string]uint32, len(res.DirPath))
	for d, p := range res.DirPath {
		if _, dup := idx[p]; !dup {
			idx[p] = uint32(d)
		}
	}
	df := &DockerfileGPUContainerizer{gpu: gpuScriptContainerizer{res, idx, m2kdetect.DockerfileOptionNames, "dockerfiles", 0}}
	s2i := &S2IGPUContainerizer{gpu: gpuScriptContainerizer{res, idx, m2kdetect.S2IOptionNames, "s2i", 12}}
	for e, feat := range res.EntryFeat { // entries are in (directory, name) order; Walk order == that order per directory
		if feat>>m2kdetect.FeatUserDF&1 == 1 {
			df.dfcontainerizers = append(df.dfcontainerizers, filepath.Dir(res.EntryPath[e]))
		}
		if feat>>m2kdetect.FeatUserS2I&1 == 1 {
			s2i.s2icontainerizers = append(s2i.s2icontainerizers, filepath.Dir(res.EntryPath[e]))
		}
	}
	// EntryFeat is in (directory id, name) order; GetFilesByName reports filepath.Walk order, which
	// interleaves a directory's files with its sub-directories by name: order by path components
	walkLess := func(a, b string) bool {
		as, bs := strings.Split(a, string(filepath.Separator)), strings.Split(b, string(filepath.Separator))
		for i := 0; i < len(as) && i < len(bs); i++ {
			if as[i] != bs[i] {
				return as[i] < bs[i]
			}
		}
		return len(as) < len(bs)
	}
	sort.SliceStable(df.dfcontainerizers, func(i, j int) bool { return walkLess(df.dfcontainerizers[i], df.dfcontainerizers[j]) })
	sort.SliceStable(s2i.s2icontainerizers, func(i, j int) bool { return walkLess(s2i.s2icontainerizers[i], s2i.s2icontainerizers[j]) })
	df.DockerfileContainerizer.Init(common.AssetsPath)
	s2i.S2IContainerizer.Init(common.AssetsPath)
	cnb := new(CNBContainerizer)
	reuse := new(ReuseContainerizer)
	c.containerizers = []Containerizer{df, s2i, cnb, reuse}
	c.rememberOutputs(res)
}

// planTimeOutputs: stdout of the accepting built-in detect scripts, keyed by option directory and service
// directory, as computed by the detection pass (m2kdetect.Result.OptionJSON).  GetContainer re-runs the detect
// script only to read this JSON (dockerfilecontainerizer.go:105-129, s2icontainerizer.go:97-115); with the two-line
// change to detect() shown in INTEGRATION.md it takes it from here instead: no fork/exec, no second grep -R.
var planTimeOutputs = map[[2]string]string{}

// CachedDetectOutput returns the plan-time stdout of scriptDir's detect script for directory, if the GPU pass has it.
func CachedDetectOutput(scriptDir, directory string) (string, bool) {
	out, ok := planTimeOutputs[[2]string{scriptDir, directory}]
	return out, ok
}

func (c *Containerizers) rememberOutputs(res *m2kdetect.Result) {
	for _, cz := range c.containerizers {
		var g *gpuScriptContainerizer
		var dirs []string
		switch v := cz.(type) {
		case *DockerfileGPUContainerizer:
			g, dirs = &v.gpu, v.dfcontainerizers
		case *S2IGPUContainerizer:
			g, dirs = &v.gpu, v.s2icontainerizers
		default:
			continue
		}
		for _, optionDir := range dirs {
			i := g.builtinIndex(optionDir)
			if i < 0 {
				continue
			}
			for d, byBit := range res.OptionJSON {
				if out, ok := byBit[int(g.shift)+i]; ok {
					planTimeOutputs[[2]string{optionDir, res.DirPath[d]}] = out
				}
			}
		}
	}
}

// NeedsHostOptions reports whether any option must be evaluated on the host per visited directory
// (user-supplied detect scripts; CNB whenever a container runtime is available).
func (c *Containerizers) NeedsHostOptions() bool {
	for _, cz := range c.containerizers {
		switch v := cz.(type) {
		case *DockerfileGPUContainerizer:
			for _, o := range v.dfcontainerizers {
				if v.gpu.builtinIndex(o) < 0 {
					return true
				}
			}
		case *S2IGPUContainerizer:
			for _, o := range v.s2icontainerizers {
				if v.gpu.builtinIndex(o) < 0 {
					return true
				}
			}
		case *CNBContainerizer:
			return true // IsBuilderSupported decides per directory (cnbcontainerizer.go:48-63)
		}
	}
	return false
}
