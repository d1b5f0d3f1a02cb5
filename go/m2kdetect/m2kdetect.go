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
	// OptionJSON[d][bit]: what the accepting built-in detect script `bit` (DirOpts bit number) prints on stdout for
	// service directory d -- the JSON GetContainer re-runs the script for at translate time
	// (dockerfilecontainerizer.go:105-129, s2icontainerizer.go:97-115); from m2k_option_json, nothing executed
	OptionJSON map[uint32]map[int]string
}

// Detect walks rootPath once and runs the whole path on the GPU.  It uses m2k_pack_detect: the file bytes
// are streamed from the files through the context's pinned blocks and never gathered on the host (the
// measured fastest road for a tree that is looked at once; page-locking a whole-tree buffer costs more
// than reading the files).  A GPU failure is returned as an error: there is no CPU fallback.
// A root that Lstat says is not a directory returns ErrNotDir (filepath.Walk visits it alone).
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

	var params C.m2k_detect_params
	if dashSh {
		params.sh_flavor = C.M2K_SH_DASH
	}
	// pre paths live in C memory for the duration of the call: never hand Go-GC memory to C code that keeps it
	n := len(preContainerized)
	carr := (**C.char)(C.malloc(C.size_t(n+1) * C.size_t(unsafe.Sizeof(uintptr(0)))))
	defer C.free(unsafe.Pointer(carr))
	cslots := (*[1 << 28]*C.char)(unsafe.Pointer(carr))[: n+1 : n+1] // Go 1.15: no unsafe.Slice
	for i, p := range preContainerized {
		cslots[i] = C.CString(p)
		defer C.free(unsafe.Pointer(cslots[i]))
	}
	cslots[n] = nil
	params.n_pre = C.uint32_t(n)
	params.pre_paths = carr

	var tree *C.m2k_tree
	var res *C.m2k_result
	if rc := C.m2k_pack_detect(c, croot, &params, &tree, &res); rc != C.M2K_OK {
		if rc == C.M2K_ERR_NOTDIR {
			return nil, ErrNotDir
		}
		return nil, fmt.Errorf("m2kdetect: detect %q: %s", rootPath, C.GoString(C.m2k_last_error(c)))
	}
	defer C.m2k_tree_free(c, tree)
	defer C.m2k_result_free(res)
	return copyOut(tree, res), nil
}

// DetectMulti is Detect on several GPUs of the box from this one process (the reference's caller is a single
// process, any2kube.go:43-46): one context per device, the tree packed once into host memory, then
// m2k_detect_multi -- a host thread per device uploads and scans a shard, the shards reduce into device 0's
// arrays over NVLink, device 0 assembles the result.  devices == nil uses context() alone.
func DetectMulti(rootPath string, preContainerized []string, dashSh bool, devices []int) (*Result, error) {
	if len(devices) <= 1 {
		return Detect(rootPath, preContainerized, dashSh)
	}
	mu.Lock()
	defer mu.Unlock()
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	n := len(devices)
	cctxs := (**C.m2k_ctx)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(uintptr(0)))))
	defer C.free(unsafe.Pointer(cctxs))
	slots := (*[1 << 20]*C.m2k_ctx)(unsafe.Pointer(cctxs))[:n:n]
	for i, d := range devices {
		var c *C.m2k_ctx
		if rc := C.m2k_ctx_create(C.int(d), &c); rc != C.M2K_OK {
			for j := 0; j < i; j++ {
				C.m2k_ctx_destroy(slots[j])
			}
			return nil, fmt.Errorf("m2kdetect: device %d: %s", d, C.GoString(C.m2k_global_error()))
		}
		slots[i] = c
	}
	defer func() {
		for _, c := range slots {
			C.m2k_ctx_destroy(c)
		}
	}()
	croot := C.CString(rootPath)
	defer C.free(unsafe.Pointer(croot))
	var tree *C.m2k_tree
	// ordinary (pageable) host memory: page-locking a whole tree costs more than reading it; the uploads stage it
	if rc := C.m2k_pack_tree(nil, croot, 1, &tree); rc != C.M2K_OK {
		if rc == C.M2K_ERR_NOTDIR {
			return nil, ErrNotDir
		}
		return nil, fmt.Errorf("m2kdetect: pack %q: %s", rootPath, C.GoString(C.m2k_global_error()))
	}
	defer C.m2k_tree_free(nil, tree)
	var params C.m2k_detect_params
	if dashSh {
		params.sh_flavor = C.M2K_SH_DASH
	}
	np := len(preContainerized)
	carr := (**C.char)(C.malloc(C.size_t(np+1) * C.size_t(unsafe.Sizeof(uintptr(0)))))
	defer C.free(unsafe.Pointer(carr))
	cslots := (*[1 << 28]*C.char)(unsafe.Pointer(carr))[: np+1 : np+1]
	for i, p := range preContainerized {
		cslots[i] = C.CString(p)
		defer C.free(unsafe.Pointer(cslots[i]))
	}
	cslots[np] = nil
	params.n_pre = C.uint32_t(np)
	params.pre_paths = carr
	var res *C.m2k_result
	if rc := C.m2k_detect_multi(cctxs, C.uint32_t(n), tree, &params, &res); rc != C.M2K_OK {
		return nil, fmt.Errorf("m2kdetect: detect %q on %d devices: %s", rootPath, n, C.GoString(C.m2k_last_error(slots[0])))
	}
	defer C.m2k_result_free(res)
	return copyOut(tree, res), nil
}

// ErrNotDir: the root exists but is not a directory for Lstat (a file, or a symlink even to a directory).
var ErrNotDir = fmt.Errorf("m2kdetect: root is not a directory")

// IsDockerFile is isDockerFile (dockerfile2kube.go:117-144) for one file's bytes, through the same kernels:
// what getDockerfileServices needs when its root is not a directory.
func IsDockerFile(data []byte) (bool, error) {
	mu.Lock()
	defer mu.Unlock()
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	c, err := context()
	if err != nil {
		return false, err
	}
	var cdata unsafe.Pointer
	if len(data) > 0 {
		cdata = C.CBytes(data)
		defer C.free(cdata)
	}
	var flags C.uint32_t
	if rc := C.m2k_is_dockerfile(c, (*C.uint8_t)(cdata), C.uint64_t(len(data)), &flags); rc != C.M2K_OK {
		return false, fmt.Errorf("m2kdetect: %s", C.GoString(C.m2k_last_error(c)))
	}
	return flags&C.M2K_FILE_DOCKERFILE != 0, nil
}

// Go 1.15-compatible views of C arrays (the reference's go.mod says go 1.15; unsafe.Slice is 1.17)
func u32s(p *C.uint32_t, n int) []uint32 {
	if n == 0 || p == nil {
		return nil
	}
	return (*[1 << 28]uint32)(unsafe.Pointer(p))[:n:n]
}
func u8s(p *C.uint8_t, n int) []uint8 {
	if n == 0 || p == nil {
		return nil
	}
	return (*[1 << 30]uint8)(unsafe.Pointer(p))[:n:n]
}
func strs(p **C.char, n int) []*C.char {
	if n == 0 || p == nil {
		return nil
	}
	return (*[1 << 28]*C.char)(unsafe.Pointer(p))[:n:n]
}

func copyOut(t *C.m2k_tree, r *C.m2k_result) *Result {
	nd, ne, nf := int(t.n_dirs), int(t.n_entries), int(t.n_files)
	out := &Result{}
	names := u8s(t.names, int(t.n_name_bytes))
	nameOff := u32s(t.entry_name_off, ne+1)
	entryDir := u32s(t.entry_dir, ne)
	dirEntry := u32s(t.dir_entry, nd)
	dirParent := u32s(t.dir_parent, nd)
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
	out.DirOpts = append([]uint32(nil), u32s(r.dir_opts, nd)...)
	out.DirState = append([]uint8(nil), u8s(r.dir_state, nd)...)
	out.DirIgnore = append([]uint8(nil), u8s(r.dir_ignore, nd)...)
	out.DirEnd = append([]uint32(nil), u32s(t.dir_end, nd)...)
	out.DirMainFile = append([]uint32(nil), u32s(r.dir_main_file, nd)...)
	out.ServiceDirs = append([]uint32(nil), u32s(r.service_dirs, int(r.n_services))...)
	out.FileFlags = append([]uint8(nil), u8s(r.file_flags, nf)...)
	out.EntryFeat = append([]uint32(nil), u32s(r.entry_feat, ne)...)
	fileEntry := u32s(t.file_entry, nf)
	out.FilePath = make([]string, nf)
	for f := 0; f < nf; f++ {
		out.FilePath[f] = out.EntryPath[fileEntry[f]]
	}
	for _, p := range strs(r.ignore_dirs, int(r.n_ignore_dirs)) {
		out.IgnoreDirs = append(out.IgnoreDirs, C.GoString(p))
	}
	for _, p := range strs(r.ignore_contents, int(r.n_ignore_contents)) {
		out.IgnoreConts = append(out.IgnoreConts, C.GoString(p))
	}
	out.KernelMillis = float32(r.ms_kernels)
	out.OptionJSON = map[uint32]map[int]string{}
	buf := (*C.char)(C.malloc(8192))
	defer C.free(unsafe.Pointer(buf))
	for _, d := range out.ServiceDirs {
		m := map[int]string{}
		for bit := 0; bit < int(C.M2K_N_OPTS); bit++ {
			if out.DirOpts[d]>>uint(bit)&1 == 0 {
				continue
			}
			if n := C.m2k_option_json(t, r, C.uint32_t(d), C.uint32_t(bit), buf, 8192); n >= 0 {
				m[bit] = C.GoStringN(buf, n)
			}
		}
		out.OptionJSON[d] = m
	}
	return out
}

// joinPath is what filepath.Walk does for every child (path/filepath.Join).
func joinPath(dir, name string) string { return filepath.Join(dir, name) }
