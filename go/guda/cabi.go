// Package guda is the drop-in host layer: the same exported names, argument meaning and error
// behaviour as the root package of LynnColeArt/guda (memory.go, guda.go, math.go, fused.go,
// fused_gelu.go, bfloat16.go, float16.go), with every hot-path call crossing cgo into
// libguda_b200.so (include/guda_b200.h) instead of running on the CPU.
//
// Build (on a machine with Go, the CUDA runtime and the built library):
//
//	CGO_CFLAGS="-I$REPO/include" CGO_LDFLAGS="-L$REPO/guda_b200/lib -lguda_b200 -Wl,-rpath,$REPO/guda_b200/lib" go test ./guda
//
// The build image of this repository has no Go toolchain, so these files are exercised there only
// through the identical C ABI from Python (guda_b200/capi.py) and C++ (host/guda.hpp).
package guda

/*
#cgo LDFLAGS: -lguda_b200
#include <stdlib.h>
#include "guda_b200.h"
*/
import "C"

import (
	"strings"
)

// status converts a C status into the reference's error model (errors.go:27-44).  BLAS argument
// violations come back as GUDA_ERR_INVALID_ARG with the exact Go panic string of
// compute/errors.go:9-35 ("blas: m < 0", ...): those are re-raised as panics, as the reference does.
func status(rc C.int, op string) error {
	if rc == C.GUDA_OK {
		return nil
	}
	msg := C.GoString(C.guda_last_error())
	if rc == C.GUDA_ERR_INVALID_ARG && strings.HasPrefix(msg, "blas: ") {
		panic(msg)
	}
	if rc == C.GUDA_ERR_NOT_IMPLEMENTED && msg == "operation not supported" {
		return ErrNotSupported
	}
	if rc == C.GUDA_ERR_MEMORY && msg == "double free detected" {
		return ErrDoubleFree
	}
	return &GUDAError{Type: ErrorType(int(rc) - 1), Op: op, Message: msg}
}

func cbool(b bool) C.int {
	if b {
		return 1
	}
	return 0
}

func (s *Stream) handle() C.guda_stream_t {
	if s == nil {
		return nil
	}
	return s.h
}
