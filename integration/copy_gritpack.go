//go:build cgo && gritpack

// Package copy, gritpack flavour: drop this file next to pkg/gritagent/copy/copy.go of the reference
// (github.com/kaito-project/grit) and build the agent with `CGO_ENABLED=1 -tags gritpack`.
// It keeps the reference's function names and error behaviour; see INTEGRATION.md.
//
// NOT COMPILED in this repository's build environment (no Go toolchain): every C entry point used here is
// exercised through the same ABI by tests/ (ctypes).
package copy

/*
#cgo CFLAGS: -I${SRCDIR}/../../../third_party/gritpack/include
#cgo LDFLAGS: -L${SRCDIR}/../../../third_party/gritpack/lib -lgritpack -lstdc++ -lpthread -ldl -lrt
#include <stdlib.h>
#include "gritpack.h"
*/
import "C"

import (
	"context"
	"errors"
	"os"
	"path/filepath"
	"unsafe"

	"sigs.k8s.io/controller-runtime/pkg/log"
)

// Direction makes explicit what the reference leaves to the Job's swapped --src-dir/--dst-dir
// (pkg/gritmanager/agentmanager/manager.go:124-134): the packed format is not symmetric.
type Direction int

const (
	Upload   Direction = iota // checkpoint.go:20: host dir -> PV
	Download                  // restore.go:16:    PV -> host dir
)

func transfer(ctx context.Context, dir Direction, srcDir, dstDir string) error {
	csrc, cdst := C.CString(srcDir), C.CString(dstDir)
	defer C.free(unsafe.Pointer(csrc))
	defer C.free(unsafe.Pointer(cdst))
	var opts C.grit_opts
	opts.struct_size = C.uint32_t(unsafe.Sizeof(opts))
	opts.device = 0
	opts.verify = 1
	if p := os.Getenv("GRIT_PARENT_DIR"); p != "" && dir == Upload { // previous checkpoint on the PV
		cp := C.CString(p)
		defer C.free(unsafe.Pointer(cp))
		opts.parent_dir = cp
	}
	var st C.grit_stats
	errbuf := (*C.char)(C.malloc(1024))
	defer C.free(unsafe.Pointer(errbuf))
	var rc C.int
	if dir == Upload {
		rc = C.grit_pack_dir(csrc, cdst, &opts, &st, errbuf, 1024)
	} else {
		rc = C.grit_restore_dir(csrc, cdst, &opts, &st, errbuf, 1024)
	}
	if rc != 0 {
		return errors.New(C.GoString(errbuf))
	}
	log.FromContext(ctx).Info("data transfer completed", "src-dir", srcDir, "dst-dir", dstDir,
		"bytes", uint64(st.bytes_logical), "stored", uint64(st.bytes_stored), "seconds", float64(st.seconds),
		"zero", uint64(st.n_zero), "dup", uint64(st.n_dup), "parent", uint64(st.n_parent))
	return nil
}

// TransferDataUpload replaces copy.TransferData at checkpoint.go:20.
func TransferDataUpload(ctx context.Context, srcDir, dstDir string) error {
	return transfer(ctx, Upload, srcDir, dstDir)
}

// TransferDataDownload replaces copy.TransferData at restore.go:16.  A checkpoint that the stock agent
// uploaded is a plain mirror without a manifest; grit_restore_dir refuses it (GRIT_EFORMAT), so it goes
// through the stock TransferData of copy.go, which stays in this package.
func TransferDataDownload(ctx context.Context, srcDir, dstDir string) error {
	if _, err := os.Stat(filepath.Join(srcDir, "gritpack.manifest")); os.IsNotExist(err) {
		return TransferData(ctx, srcDir, dstDir)
	}
	return transfer(ctx, Download, srcDir, dstDir)
}

// CreateSentinelFile keeps the reference signature and payload (copy.go:92-102).
func CreateSentinelFile(dir, fileName string) error {
	cd, cf := C.CString(dir), C.CString(fileName)
	defer C.free(unsafe.Pointer(cd))
	defer C.free(unsafe.Pointer(cf))
	if C.grit_write_sentinel(cd, cf) != 0 {
		return errors.New("create sentinel file " + dir + "/" + fileName + " failed")
	}
	return nil
}

// Warmup brings the CUDA runtime up on a goroutine of its own.  A process's first CUDA call costs 0.3-0.5 s whatever it
// does: RunCheckpoint calls Warmup before RuntimeCheckpointPod (pkg/gritagent/checkpoint/checkpoint.go:14), so that the
// cost hides behind the CRIU dump, and RunRestore calls it first thing.
func Warmup() {
	go func() { C.grit_warmup(0) }()
}
