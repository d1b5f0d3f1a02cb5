// Package m2kdetect is the cgo binding of libm2kdetect (include/m2kdetect.h): GPU source detection
// for move2kube's planning pass.
//
// NOT COMPILED IN THIS REPOSITORY'S CI: neither the build container nor the GPU box has a Go
// toolchain (`go: command not found`), so this file is the binding a maintainer would add next to
// internal/source and internal/containerizer, written against the C ABI that IS tested (through
// ctypes).  It keeps the reference's interfaces: see any2kube_gpu.go for the Translator and
// containerizers_gpu.go for the Containerizer implementations that call into it.
package m2kdetect

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../move2kube_b200 -lm2kdetect -Wl,-rpath,${SRCDIR}/../../move2kube_b200
#include <stdlib.h>
#include "m2kdetect.h"
*/
import "C"

import (
	"fmt"
	"path/filepath"
	"runtime"
	"sync"
	"unsafe"
)

// Option bit numbers of Result.DirOpts (TargetOptions order of the reference, SURVEY.md App. A).
var DockerfileOptionNames = []string{"django", "golang", "java-war-jboss", "java-war-liberty", "java-war-tomcat",
	"javaant", "javagradle", "javamaven", "nodejs", "php", "python", "ruby"}
var S2IOptionNames = []string{"golang", "java", "nodejs", "php", "python", "ruby"}

// Directory states of the walk (any2kube.go:58-98).
const (
	DirNotVisited = iota
	DirPrecontainerized
	DirIgnoredPrune
	DirIgnoredDescend
	DirNomatchPrune
	DirNomatchDescend
	DirService
	DirWalkError
)

// Name-feature bits of Result.EntryFeat that host code asks for.
const (
	FeatIgnoreFile = 14
	FeatUserDF     = 15
	FeatUserS2I    = 16
	FeatYAML       = 17
)

var (
	ctxOnce sync.Once
	ctx     *C.m2k_ctx
	ctxErr  error
	mu      sync.Mutex // the reference path is single-goroutine (SURVEY.md section 5); so is a context
)

// context returns the process-wide device context.  The reference builds a new Containerizers per
// call (any2kube.go:45, cfmanifest2kube.go:53); the GPU context must outlive them.
func context() (*C.m2k_ctx, error) {
	ctxOnce.Do(func() {
		var c *C.m2k_ctx
		if rc := C.m2k_ctx_create(0, &c); rc != C.M2K_OK {
			ctxErr = fmt.Errorf("m2kdetect: %s", C.GoString(C.m2k_global_error()))
			return
		}
		ctx = c
	})
	return ctx, ctxErr
}

// Result is one detection pass over a tree, copied into Go memory.
type Result struct {
	DirPath      []string // filepath.Walk path of every directory, pre-order
	DirOpts      []uint32 // built-in option bits per directory
	DirState     []uint8
	DirIgnore    []uint8  // bit0 ignoreDirectories, bit1 ignoreContents, bit2 preContainerized
	DirEnd       []uint32 // subtree end (pre-order ids), for SkipDir
	DirMainFile  []uint32
	ServiceDirs  []uint32
	FileFlags    []uint8
	FilePath     []string
	EntryFeat    []uint32
	EntryPath    []string
	IgnoreDirs   []string
	IgnoreConts  []string
	KernelMillis float32
}

// Detect packs rootPath (one walk, pinned host buffers owned by C) and runs the whole path on the GPU.
// A GPU failure is returned as an error: there is no CPU fallback.
func Detect(rootPath string, preContainerized []string, dashSh bool) (*Result, error) {
	mu.Lock()
	defer mu.Unlock()
	runtime.LockOSThread() // CUDA's current device is per OS thread
	defer runtime.UnlockOSThread()
	c, err := context()
	if err != nil {
		return nil, err
	}
	croot := C.CString(rootPath)
	defer C.free(unsafe.Pointer(croot))
	var tree *C.m2k_tree
	if rc := C.m2k_pack_tree(c, croot, 1, &tree); rc != C.M2K_OK {
		return nil, fmt.Errorf("m2kdetect: pack %q: %s", rootPath, C.GoString(C.m2k_last_error(c)))
	}
	defer C.m2k_tree_free(c, tree)

	var params C.m2k_detect_params
	if dashSh {
		params.sh_flavor = C.M2K_SH_DASH
	}
	// pre paths live in C memory for the duration of the call: never hand Go-GC memory to async copies
	cpre := make([]*C.char, len(preContainerized)+1)
	for i, p := range preContainerized {
		cpre[i] = C.CString(p)
		defer C.free(unsafe.Pointer(cpre[i]))
	}
	carr := (**C.char)(C.malloc(C.size_t(len(cpre)) * C.size_t(unsafe.Sizeof(uintptr(0)))))
	defer C.free(unsafe.Pointer(carr))
	copy(unsafe.Slice(carr, len(cpre)), cpre)
	params.n_pre = C.uint32_t(len(preContainerized))
	params.pre_paths = carr

	var res *C.m2k_result
	if rc := C.m2k_detect(c, tree, &params, &res); rc != C.M2K_OK {
		return nil, fmt.Errorf("m2kdetect: detect %q: %s", rootPath, C.GoString(C.m2k_last_error(c)))
	}
	defer C.m2k_result_free(res)
	return copyOut(tree, res), nil
}

func copyOut(t *C.m2k_tree, r *C.m2k_result) *Result {
	nd, ne, nf := int(t.n_dirs), int(t.n_entries), int(t.n_files)
	out := &Result{}
	names := unsafe.Slice((*byte)(unsafe.Pointer(t.names)), int(t.n_name_bytes))
	nameOff := unsafe.Slice((*uint32)(unsafe.Pointer(t.entry_name_off)), ne+1)
	entryDir := unsafe.Slice((*uint32)(unsafe.Pointer(t.entry_dir)), ne)
	dirEntry := unsafe.Slice((*uint32)(unsafe.Pointer(t.dir_entry)), nd)
	dirParent := unsafe.Slice((*uint32)(unsafe.Pointer(t.dir_parent)), nd)
	root := C.GoString(t.root_path)
	out.DirPath = make([]string, nd)
	for d := 0; d < nd; d++ {
		if d == 0 {
			out.DirPath[0] = root // Walk reports the root as given
			continue
		}
		e := dirEntry[d]
		out.DirPath[d] = joinPath(out.DirPath[dirParent[d]], string(names[nameOff[e]:nameOff[e+1]]))
	}
	out.EntryPath = make([]string, ne)
	for e := 0; e < ne; e++ {
		out.EntryPath[e] = joinPath(out.DirPath[entryDir[e]], string(names[nameOff[e]:nameOff[e+1]]))
	}
	out.DirOpts = append([]uint32(nil), unsafe.Slice((*uint32)(unsafe.Pointer(r.dir_opts)), nd)...)
	out.DirState = append([]uint8(nil), unsafe.Slice((*uint8)(unsafe.Pointer(r.dir_state)), nd)...)
	out.DirIgnore = append([]uint8(nil), unsafe.Slice((*uint8)(unsafe.Pointer(r.dir_ignore)), nd)...)
	out.DirEnd = append([]uint32(nil), unsafe.Slice((*uint32)(unsafe.Pointer(t.dir_end)), nd)...)
	out.DirMainFile = append([]uint32(nil), unsafe.Slice((*uint32)(unsafe.Pointer(r.dir_main_file)), nd)...)
	out.ServiceDirs = append([]uint32(nil), unsafe.Slice((*uint32)(unsafe.Pointer(r.service_dirs)), int(r.n_services))...)
	out.FileFlags = append([]uint8(nil), unsafe.Slice((*uint8)(unsafe.Pointer(r.file_flags)), nf)...)
	out.EntryFeat = append([]uint32(nil), unsafe.Slice((*uint32)(unsafe.Pointer(r.entry_feat)), ne)...)
	fileEntry := unsafe.Slice((*uint32)(unsafe.Pointer(t.file_entry)), nf)
	out.FilePath = make([]string, nf)
	for f := 0; f < nf; f++ {
		out.FilePath[f] = out.EntryPath[fileEntry[f]]
	}
	for _, p := range unsafe.Slice(r.ignore_dirs, int(r.n_ignore_dirs)) {
		out.IgnoreDirs = append(out.IgnoreDirs, C.GoString(p))
	}
	for _, p := range unsafe.Slice(r.ignore_contents, int(r.n_ignore_contents)) {
		out.IgnoreConts = append(out.IgnoreConts, C.GoString(p))
	}
	out.KernelMillis = float32(r.ms_kernels)
	return out
}

// joinPath is what filepath.Walk does for every child (path/filepath.Join).
func joinPath(dir, name string) string { return filepath.Join(dir, name) }
