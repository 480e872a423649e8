package guda

/*
#include "guda_b200.h"
*/
import "C"

// Sgemv mirrors compute/level2float32.go:244-308 on device memory (row-major, lda = row stride):
//
//	y = alpha * A * x + beta * y   (trans == false)
//	y = alpha * Aᵀ * x + beta * y  (trans == true)
//
// The panics come in the reference's order: the dimension / leading-dimension / increment checks are repeated by the
// C library with the same strings (status() re-raises them), the slice-length checks are made here because they need
// the DevicePtr sizes.  Results are bit-identical to the reference's AVX2 route for every increment sign.
func Sgemv(trans bool, m, n int, alpha float32, a DevicePtr, lda int, x DevicePtr, incX int, beta float32, y DevicePtr, incY int) error {
	if m >= 0 && n >= 0 && lda >= imax(1, n) && incX != 0 && incY != 0 && m > 0 && n > 0 {
		lenX, lenY := n, m
		if trans {
			lenX, lenY = m, n
		}
		lx, ly := x.size/4, y.size/4
		if (incX > 0 && (lenX-1)*incX >= lx) || (incX < 0 && (1-lenX)*incX >= lx) {
			panic(shortX)
		}
		if (incY > 0 && (lenY-1)*incY >= ly) || (incY < 0 && (1-lenY)*incY >= ly) {
			panic(shortY)
		}
		if a.size/4 < lda*(m-1)+n {
			panic(shortA)
		}
	}
	if err := status(C.guda_sgemv(cbool(trans), C.int(m), C.int(n), C.float(alpha), (*C.float)(a.ptr), C.int(lda),
		(*C.float)(x.ptr), C.int(incX), C.float(beta), (*C.float)(y.ptr), C.int(incY), nil), "Sgemv"); err != nil {
		return err
	}
	// compute/asm/f32/gemv.go:61-64: with beta == 0 and incY == 1 GemvT clears the whole y slice, not only y[:n]
	if trans && m > 0 && n > 0 && alpha != 0 && beta == 0 && incY == 1 && y.size/4 > n {
		tail := y.Offset(4 * n)
		return status(C.guda_memset(tail.ptr, 0, C.size_t(tail.size/4*4), nil), "Sgemv")
	}
	return nil
}

// Sger mirrors compute/level2float32.go:24-65: A += alpha * x * yᵀ (row-major), every element computed as
// (alpha*x_i)*y_j rounded, then added to a_ij (compute/asm/f32/ge_amd64.s).
func Sger(m, n int, alpha float32, x DevicePtr, incX int, y DevicePtr, incY int, a DevicePtr, lda int) error {
	if m >= 0 && n >= 0 && lda >= imax(1, n) && incX != 0 && incY != 0 && m > 0 && n > 0 {
		lx, ly := x.size/4, y.size/4
		if (incX > 0 && lx <= (m-1)*incX) || (incX < 0 && lx <= (1-m)*incX) {
			panic(shortX)
		}
		if (incY > 0 && ly <= (n-1)*incY) || (incY < 0 && ly <= (1-n)*incY) {
			panic(shortY)
		}
		if a.size/4 < lda*(m-1)+n {
			panic(shortA)
		}
	}
	return status(C.guda_sger(C.int(m), C.int(n), C.float(alpha), (*C.float)(x.ptr), C.int(incX), (*C.float)(y.ptr), C.int(incY),
		(*C.float)(a.ptr), C.int(lda), nil), "Sger")
}

// Dgemv mirrors compute/level2float64.go:232-300 (fp64 twin of Sgemv; kernels compute/asm/f64/gemvN_avx2_amd64.s and
// gemvT_amd64.s).  Bit-identical to the reference's AVX2+FMA route for every increment sign.  Unlike the fp32 GemvT, the
// fp64 one clears or scales exactly n elements of y (compute/asm/f64/gemvT_amd64.s), so there is no whole-slice clear to reproduce.
func Dgemv(trans bool, m, n int, alpha float64, a DevicePtr, lda int, x DevicePtr, incX int, beta float64, y DevicePtr, incY int) error {
	if m >= 0 && n >= 0 && lda >= imax(1, n) && incX != 0 && incY != 0 && m > 0 && n > 0 {
		lenX, lenY := n, m
		if trans {
			lenX, lenY = m, n
		}
		lx, ly := x.size/8, y.size/8
		if (incX > 0 && (lenX-1)*incX >= lx) || (incX < 0 && (1-lenX)*incX >= lx) {
			panic(shortX)
		}
		if (incY > 0 && (lenY-1)*incY >= ly) || (incY < 0 && (1-lenY)*incY >= ly) {
			panic(shortY)
		}
		if a.size/8 < lda*(m-1)+n {
			panic(shortA)
		}
	}
	return status(C.guda_dgemv(cbool(trans), C.int(m), C.int(n), C.double(alpha), (*C.double)(a.ptr), C.int(lda),
		(*C.double)(x.ptr), C.int(incX), C.double(beta), (*C.double)(y.ptr), C.int(incY), nil), "Dgemv")
}

// Dger mirrors compute/level2float64.go:24-65: A += alpha * x * yᵀ on float64 (compute/asm/f64/ger_amd64.s).
func Dger(m, n int, alpha float64, x DevicePtr, incX int, y DevicePtr, incY int, a DevicePtr, lda int) error {
	if m >= 0 && n >= 0 && lda >= imax(1, n) && incX != 0 && incY != 0 && m > 0 && n > 0 {
		lx, ly := x.size/8, y.size/8
		if (incX > 0 && lx <= (m-1)*incX) || (incX < 0 && lx <= (1-m)*incX) {
			panic(shortX)
		}
		if (incY > 0 && ly <= (n-1)*incY) || (incY < 0 && ly <= (1-n)*incY) {
			panic(shortY)
		}
		if a.size/8 < lda*(m-1)+n {
			panic(shortA)
		}
	}
	return status(C.guda_dger(C.int(m), C.int(n), C.double(alpha), (*C.double)(x.ptr), C.int(incX), (*C.double)(y.ptr), C.int(incY),
		(*C.double)(a.ptr), C.int(lda), nil), "Dger")
}
