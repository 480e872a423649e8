package guda

/*
#include "guda_b200.h"
*/
import "C"

// Panic strings of compute/errors.go:9-35.  GEMM's checks are made here, in the reference's order, before crossing
// cgo (the slice-length checks need the DevicePtr sizes); the C library repeats the dimension / leading-dimension /
// increment checks and returns the same strings, which status() re-raises as panics.
const (
	shortX = "blas: insufficient length of x"
	shortY = "blas: insufficient length of y"
	shortA = "blas: insufficient length of a"
	shortB = "blas: insufficient length of b"
	shortC = "blas: insufficient length of c"
)

func imax(a, b int) int {
	if a > b {
		return a
	}
	return b
}

// AXPY mirrors math.go:39-47: y = alpha*x + y over n float32 elements (one FMA per element, bit-exact
// with the reference's AVX2+FMA route).  Stream-ordered and asynchronous on the default stream.
func AXPY(alpha float32, x, y DevicePtr, n int) error {
	if n > x.size/4 || n > y.size/4 {
		panic("runtime error: slice bounds out of range") // x.Float32()[:n] in the reference
	}
	return status(C.guda_saxpy(C.int(n), C.float(alpha), (*C.float)(x.ptr), 1, (*C.float)(y.ptr), 1, nil), "AXPY")
}

// DOT mirrors math.go:64-71; blocks until the scalar is on the host.
func DOT(x, y DevicePtr, n int) (float32, error) {
	if n > x.size/4 || n > y.size/4 {
		panic("runtime error: slice bounds out of range")
	}
	var r C.float
	err := status(C.guda_sdot(C.int(n), (*C.float)(x.ptr), 1, (*C.float)(y.ptr), 1, &r, nil), "DOT")
	return float32(r), err
}

// Scale mirrors math.go:170-177.
func Scale(alpha float32, x DevicePtr, n int) error {
	if n > x.size/4 {
		panic("runtime error: slice bounds out of range")
	}
	return status(C.guda_sscal(C.int(n), C.float(alpha), (*C.float)(x.ptr), 1, nil), "Scale")
}

// checkGemm makes the reference's argument checks in the reference's order (compute/sgemm.go:30-76,116-136) and
// reports whether there is anything to compute.
func checkGemm(transA, transB bool, m, n, k, lda, ldb, ldc, lenA, lenB, lenC int) bool {
	if m < 0 {
		panic("blas: m < 0")
	}
	if n < 0 {
		panic("blas: n < 0")
	}
	if k < 0 {
		panic("blas: k < 0")
	}
	if transA {
		if lda < imax(1, m) {
			panic("blas: bad leading dimension of A")
		}
	} else if lda < imax(1, k) {
		panic("blas: bad leading dimension of A")
	}
	if transB {
		if ldb < imax(1, k) {
			panic("blas: bad leading dimension of B")
		}
	} else if ldb < imax(1, n) {
		panic("blas: bad leading dimension of B")
	}
	if ldc < imax(1, n) {
		panic("blas: bad leading dimension of C")
	}
	if m == 0 || n == 0 {
		return false
	}
	if transA {
		if lenA < (k-1)*lda+m {
			panic(shortA)
		}
	} else if lenA < (m-1)*lda+k {
		panic(shortA)
	}
	if transB {
		if lenB < (n-1)*ldb+k {
			panic(shortB)
		}
	} else if lenB < (k-1)*ldb+n {
		panic(shortB)
	}
	if lenC < (m-1)*ldc+n {
		panic(shortC)
	}
	return true
}

// GEMM mirrors math.go:98-129 -> compute/sgemm.go:29: row-major C = alpha*op(A)*op(B) + beta*C.
// NN/TN/TT results are bit-identical to the reference's AVX2+FMA route; NT is within GEMMArchTolerance.
func GEMM(transA, transB bool, m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int,
	beta float32, c DevicePtr, ldc int) error {
	return gemmOn(nil, transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc)
}

func gemmOn(s *Stream, transA, transB bool, m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int,
	beta float32, c DevicePtr, ldc int) error {
	if !checkGemm(transA, transB, m, n, k, lda, ldb, ldc, a.size/4, b.size/4, c.size/4) {
		return nil
	}
	return status(C.guda_sgemm(cbool(transA), cbool(transB), C.int(m), C.int(n), C.int(k), C.float(alpha),
		(*C.float)(a.ptr), C.int(lda), (*C.float)(b.ptr), C.int(ldb), C.float(beta), (*C.float)(c.ptr), C.int(ldc), s.handle()), "GEMM")
}

// The cuBLAS-style names of docs/manual/08-blas-api.md:58,96,199.
func Saxpy(n int, alpha float32, x DevicePtr, incX int, y DevicePtr, incY int) error {
	return status(C.guda_saxpy(C.int(n), C.float(alpha), (*C.float)(x.ptr), C.int(incX), (*C.float)(y.ptr), C.int(incY), nil), "Saxpy")
}
func Sdot(n int, x DevicePtr, incX int, y DevicePtr, incY int) (float32, error) {
	var r C.float
	err := status(C.guda_sdot(C.int(n), (*C.float)(x.ptr), C.int(incX), (*C.float)(y.ptr), C.int(incY), &r, nil), "Sdot")
	return float32(r), err
}
func Sscal(n int, alpha float32, x DevicePtr, incX int) error {
	return status(C.guda_sscal(C.int(n), C.float(alpha), (*C.float)(x.ptr), C.int(incX), nil), "Sscal")
}
func Daxpy(n int, alpha float64, x DevicePtr, incX int, y DevicePtr, incY int) error {
	return status(C.guda_daxpy(C.int(n), C.double(alpha), (*C.double)(x.ptr), C.int(incX), (*C.double)(y.ptr), C.int(incY), nil), "Daxpy")
}
func Ddot(n int, x DevicePtr, incX int, y DevicePtr, incY int) (float64, error) {
	var r C.double
	err := status(C.guda_ddot(C.int(n), (*C.double)(x.ptr), C.int(incX), (*C.double)(y.ptr), C.int(incY), &r, nil), "Ddot")
	return float64(r), err
}
func Dscal(n int, alpha float64, x DevicePtr, incX int) error {
	return status(C.guda_dscal(C.int(n), C.double(alpha), (*C.double)(x.ptr), C.int(incX), nil), "Dscal")
}
func Sgemm(transA, transB bool, m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int,
	beta float32, c DevicePtr, ldc int) error {
	return GEMM(transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc)
}

// Dgemm: compute/dgemm.go:25 (the reference has no root-level wrapper).  NN/TN run on the FP64 tensor pipe.
func Dgemm(transA, transB bool, m, n, k int, alpha float64, a DevicePtr, lda int, b DevicePtr, ldb int,
	beta float64, c DevicePtr, ldc int) error {
	if !checkGemm(transA, transB, m, n, k, lda, ldb, ldc, a.size/8, b.size/8, c.size/8) {
		return nil
	}
	return status(C.guda_dgemm(cbool(transA), cbool(transB), C.int(m), C.int(n), C.int(k), C.double(alpha),
		(*C.double)(a.ptr), C.int(lda), (*C.double)(b.ptr), C.int(ldb), C.double(beta), (*C.double)(c.ptr), C.int(ldc), nil), "Dgemm")
}

// Stream-ordered variants (new): the same kernels on a caller-chosen CUDA stream.
func (s *Stream) AXPY(alpha float32, x, y DevicePtr, n int) error {
	return status(C.guda_saxpy(C.int(n), C.float(alpha), (*C.float)(x.ptr), 1, (*C.float)(y.ptr), 1, s.handle()), "AXPY")
}
func (s *Stream) GEMM(transA, transB bool, m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int,
	beta float32, c DevicePtr, ldc int) error {
	return gemmOn(s, transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc)
}

// SetSgemmMode selects the SGEMM arithmetic: 0 = exact-FP32 SIMT (bit-exact, default), 1 = 3xTF32 tensor cores.
func SetSgemmMode(mode int) error { return status(C.guda_set_sgemm_mode(C.int(mode)), "SetSgemmMode") }

// GEMMHost is GEMM for operands that live in host memory (Go slices): the uploads are pipelined under the kernel by
// K chunks and C streams back row panel by row panel, so the transfers cost one chunk in and one panel out instead of
// the whole of A, B and C.  Same arguments, panics and bits as GEMM.  Slices from MallocHost (pinned) make the copies
// asynchronous; ordinary Go slices work too (pin them for the duration of the call with runtime.Pinner).
func GEMMHost(transA, transB bool, m, n, k int, alpha float32, a []float32, lda int, b []float32, ldb int,
	beta float32, c []float32, ldc int) error {
	if !checkGemm(transA, transB, m, n, k, lda, ldb, ldc, len(a), len(b), len(c)) {
		return nil
	}
	var pa, pb *C.float
	if len(a) > 0 {
		pa = (*C.float)(&a[0])
	}
	if len(b) > 0 {
		pb = (*C.float)(&b[0])
	}
	return status(C.guda_sgemm_host(cbool(transA), cbool(transB), C.int(m), C.int(n), C.int(k), C.float(alpha),
		pa, C.int(lda), pb, C.int(ldb), C.float(beta), (*C.float)(&c[0]), C.int(ldc), nil), "GEMM")
}

// Dsdot mirrors compute/level1float32_dsdot.go:18-54: dot product of two float32 vectors accumulated in float64
// (the products are exact).  Blocks until the scalar is on the host.
func Dsdot(n int, x DevicePtr, incX int, y DevicePtr, incY int) float64 {
	if !checkDsdot(n, x.size/4, incX, y.size/4, incY) {
		return 0
	}
	var r C.double
	if err := status(C.guda_dsdot(C.int(n), (*C.float)(x.ptr), C.int(incX), (*C.float)(y.ptr), C.int(incY), &r, nil), "Dsdot"); err != nil {
		panic(err)
	}
	return float64(r)
}

// Sdsdot mirrors compute/level1float32_sdsdot.go:18-54: alpha + float32(Dsdot); 0 when n == 0.
func Sdsdot(n int, alpha float32, x DevicePtr, incX int, y DevicePtr, incY int) float32 {
	if !checkDsdot(n, x.size/4, incX, y.size/4, incY) {
		return 0
	}
	var r C.float
	if err := status(C.guda_sdsdot(C.int(n), C.float(alpha), (*C.float)(x.ptr), C.int(incX), (*C.float)(y.ptr), C.int(incY), &r, nil), "Sdsdot"); err != nil {
		panic(err)
	}
	return float32(r)
}

// checkDsdot makes the reference's checks in its order (level1float32_dsdot.go:19-52); false = nothing to compute.
func checkDsdot(n, lenX, incX, lenY, incY int) bool {
	if incX == 0 {
		panic("blas: zero x index increment")
	}
	if incY == 0 {
		panic("blas: zero y index increment")
	}
	if n <= 0 {
		if n == 0 {
			return false
		}
		panic("blas: n < 0")
	}
	if incX == 1 && incY == 1 {
		if lenX < n {
			panic(shortX)
		}
		if lenY < n {
			panic(shortY)
		}
		return true
	}
	ix, iy := 0, 0
	if incX < 0 {
		ix = (-n + 1) * incX
	}
	if incY < 0 {
		iy = (-n + 1) * incY
	}
	if ix >= lenX || ix+(n-1)*incX >= lenX {
		panic(shortX)
	}
	if iy >= lenY || iy+(n-1)*incY >= lenY {
		panic(shortY)
	}
	return true
}
