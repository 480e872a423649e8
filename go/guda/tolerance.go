package guda

/*
#include <stdlib.h>
#include "guda_b200.h"
*/
import "C"

import (
	"fmt"
	"math"
	"unsafe"
)

// The architecture-aware tolerance framework of the reference (tolerance.go:10-191, tolerance_arch.go:8-127,
// numerical_parity.go:103-141) with the "sm_100a" architecture entry.  Every number comes from the C library
// (guda_b200/csrc/tolerance.cu) so that Go, C++ and Python callers agree; the scalar predicates are evaluated here, and
// VerifyDeviceFloat32Array runs VerifyFloat32Array's loop over two device buffers in one streaming pass on the GPU.

// ToleranceConfig is tolerance.go:10-26; KScale, HalfULPTol, Abs64 and Rel64 are what the sm_100a entry adds (zero = not
// applicable): the K-scaled forward-error coefficient, the bound in ULPs of a 16-bit output format, and the fp64 bounds.
type ToleranceConfig struct {
	AbsTol     float32
	RelTol     float32
	ULPTol     int
	CheckNaN   bool
	CheckInf   bool
	KScale     float32
	HalfULPTol int
	Abs64      float64
	Rel64      float64
}

func (t ToleranceConfig) c() C.guda_tolerance_t {
	return C.guda_tolerance_t{abs_tol: C.float(t.AbsTol), rel_tol: C.float(t.RelTol), ulp_tol: C.int(t.ULPTol), check_nan: cbool(t.CheckNaN),
		check_inf: cbool(t.CheckInf), k_scale: C.float(t.KScale), half_ulp_tol: C.int(t.HalfULPTol), abs64: C.double(t.Abs64), rel64: C.double(t.Rel64)}
}

func fromC(t C.guda_tolerance_t) ToleranceConfig {
	return ToleranceConfig{AbsTol: float32(t.abs_tol), RelTol: float32(t.rel_tol), ULPTol: int(t.ulp_tol), CheckNaN: t.check_nan != 0,
		CheckInf: t.check_inf != 0, KScale: float32(t.k_scale), HalfULPTol: int(t.half_ulp_tol), Abs64: float64(t.abs64), Rel64: float64(t.rel64)}
}

func preset(which int) ToleranceConfig {
	var t C.guda_tolerance_t
	if err := status(C.guda_tolerance_preset(C.int(which), &t), "TolerancePreset"); err != nil {
		panic(err)
	}
	return fromC(t)
}

// DefaultTolerance, StrictTolerance, RelaxedTolerance are tolerance.go:28-54.
func DefaultTolerance() ToleranceConfig { return preset(0) }
func StrictTolerance() ToleranceConfig  { return preset(1) }
func RelaxedTolerance() ToleranceConfig { return preset(2) }

// Float32ULPDiff is tolerance.go:100-118: distance in representable values, MaxInt32 across a sign change.
func Float32ULPDiff(a, b float32) int {
	x, y := math.Float32bits(a), math.Float32bits(b)
	switch {
	case (x^y)>>31 != 0:
		return math.MaxInt32
	case x >= y:
		return int(x - y)
	default:
		return int(y - x)
	}
}

// Float32NearEqual is tolerance.go:57-97: equal, or inside one of the absolute / relative / ULP bounds.
func Float32NearEqual(a, b float32, tol ToleranceConfig) bool {
	a64, b64 := float64(a), float64(b)
	if tol.CheckNaN && a != a && b != b {
		return true
	}
	if tol.CheckInf && ((math.IsInf(a64, 1) && math.IsInf(b64, 1)) || (math.IsInf(a64, -1) && math.IsInf(b64, -1))) {
		return true
	}
	if a == b {
		return true
	}
	gap := math.Abs(float64(a - b)) // float32 subtraction first, as the reference
	if gap <= float64(tol.AbsTol) || gap <= math.Max(math.Abs(a64), math.Abs(b64))*float64(tol.RelTol) {
		return true
	}
	return tol.ULPTol > 0 && Float32ULPDiff(a, b) <= tol.ULPTol
}

// VerificationResult is tolerance.go:121-128.
type VerificationResult struct {
	MaxAbsError float32
	MaxRelError float32
	MaxULPError int
	NumErrors   int
	TotalItems  int
	FirstError  int
}

// VerifyFloat32Array is tolerance.go:131-174 over host slices.
func VerifyFloat32Array(expected, actual []float32, tol ToleranceConfig) VerificationResult {
	res := VerificationResult{TotalItems: len(expected), FirstError: -1}
	if len(actual) != len(expected) {
		res.NumErrors = len(expected)
		return res
	}
	for i, want := range expected {
		got := actual[i]
		if Float32NearEqual(want, got, tol) {
			continue
		}
		if res.NumErrors++; res.FirstError < 0 {
			res.FirstError = i
		}
		ad := float32(math.Abs(float64(want - got)))
		if ad > res.MaxAbsError {
			res.MaxAbsError = ad
		}
		if want != 0 {
			if rd := ad / float32(math.Abs(float64(want))); rd > res.MaxRelError {
				res.MaxRelError = rd
			}
		}
		if ud := Float32ULPDiff(want, got); ud > res.MaxULPError {
			res.MaxULPError = ud
		}
	}
	return res
}

// VerifyDeviceFloat32Array is VerifyFloat32Array for two device buffers of n float32: one HBM-streaming pass on the GPU
// (guda_verify_f32); every field has the bits the host loop would produce.
func VerifyDeviceFloat32Array(expected, actual DevicePtr, n int, tol ToleranceConfig) (VerificationResult, error) {
	ct := tol.c()
	var r C.guda_verify_result_t
	err := status(C.guda_verify_f32((*C.float)(expected.ptr), (*C.float)(actual.ptr), C.longlong(n), &ct, &r, nil), "VerifyFloat32Array")
	return VerificationResult{MaxAbsError: float32(r.max_abs_error), MaxRelError: float32(r.max_rel_error), MaxULPError: int(r.max_ulp_error),
		NumErrors: int(r.num_errors), TotalItems: int(r.total_items), FirstError: int(r.first_error)}, err
}

// IsAcceptable is tolerance.go:177-182.
func (r VerificationResult) IsAcceptable(tol ToleranceConfig) bool {
	if r.NumErrors == 0 {
		return true
	}
	return r.MaxAbsError <= tol.AbsTol && r.MaxRelError <= tol.RelTol && r.MaxULPError <= tol.ULPTol
}

// String is tolerance.go:185-200.
func (r VerificationResult) String() string {
	if r.NumErrors == 0 {
		return "PASS: All values match within tolerance"
	}
	return fmt.Sprintf("FAIL: %d/%d values differ (%.2f%%)\n  Max absolute error: %e\n  Max relative error: %e\n  Max ULP error: %d\n  First error at index: %d",
		r.NumErrors, r.TotalItems, 100*float64(r.NumErrors)/float64(r.TotalItems), r.MaxAbsError, r.MaxRelError, r.MaxULPError, r.FirstError)
}

// ArchToleranceConfig is tolerance_arch.go:8-17 plus the SM100a override.
type ArchToleranceConfig struct {
	Base    ToleranceConfig
	AMD64   *ToleranceConfig
	ARM64   *ToleranceConfig
	Generic *ToleranceConfig
	SM100a  *ToleranceConfig
}

// ComputeArch is the architecture whose rounding behaviour the results carry: the device, not runtime.GOARCH.
const ComputeArch = "sm_100a"

func mergeTolerances(base, override ToleranceConfig) ToleranceConfig { // tolerance_arch.go:41-57
	b, o := base.c(), override.c()
	var out C.guda_tolerance_t
	if err := status(C.guda_merge_tolerances(&b, &o, &out), "mergeTolerances"); err != nil {
		panic(err)
	}
	return fromC(out)
}

// GetArchTolerance is tolerance_arch.go:20-38 with the device architecture in place of runtime.GOARCH.
func GetArchTolerance(config ArchToleranceConfig) ToleranceConfig {
	if config.SM100a == nil {
		return config.Base
	}
	return mergeTolerances(config.Base, *config.SM100a)
}

func operationTolerance(operation, arch string) ToleranceConfig {
	op, ar := C.CString(operation), C.CString(arch)
	defer C.free(unsafe.Pointer(op))
	defer C.free(unsafe.Pointer(ar))
	var t C.guda_tolerance_t
	if err := status(C.guda_get_operation_tolerance(op, ar, &t), "GetOperationTolerance"); err != nil {
		panic(err)
	}
	return fromC(t)
}

func archConfig(operation string) ArchToleranceConfig {
	cfg := ArchToleranceConfig{Base: operationTolerance(operation, "amd64")}
	for _, e := range []struct {
		arch string
		dst  **ToleranceConfig
	}{{"arm64", &cfg.ARM64}, {"generic", &cfg.Generic}, {ComputeArch, &cfg.SM100a}} {
		if t := operationTolerance(operation, e.arch); t != cfg.Base {
			v := t
			*e.dst = &v
		}
	}
	return cfg
}

// GEMMArchTolerance, ConvArchTolerance, ReduceArchTolerance are tolerance_arch.go:59-113.
var (
	GEMMArchTolerance   = archConfig("gemm")
	ConvArchTolerance   = archConfig("conv2d")
	ReduceArchTolerance = archConfig("reduce_sum")
)

// GetOperationTolerance is tolerance_arch.go:116-127; besides "gemm", "conv2d", "reduce_sum", "softmax" it knows "dot", "dgemm",
// "gemm_3xtf32", "gemm_bf16", "gemm_f16", "gemm_bf16_fused"; any other name yields DefaultTolerance().
func GetOperationTolerance(operation string) ToleranceConfig { return operationTolerance(operation, ComputeArch) }

// IsARM64 is tolerance_arch.go:130-132: the compute device is never ARM64.
func IsARM64() bool { return false }

// OperationTolerance and StandardTolerances are numerical_parity.go:124-141.
type OperationTolerance struct {
	Name   string
	AbsTol float32
	RelTol float32
	ULPTol int32
}

var StandardTolerances = func() map[string]OperationTolerance {
	m := map[string]OperationTolerance{}
	for _, name := range []string{"add", "multiply", "gemm", "conv2d", "reduce_sum", "softmax", "gelu", "exp"} {
		cn := C.CString(name)
		var a, r C.float
		var u C.int
		if C.guda_standard_tolerance(cn, &a, &r, &u) == C.GUDA_OK {
			m[name] = OperationTolerance{Name: name, AbsTol: float32(a), RelTol: float32(r), ULPTol: int32(u)}
		}
		C.free(unsafe.Pointer(cn))
	}
	return m
}()
