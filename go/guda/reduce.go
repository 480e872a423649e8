package guda

/*
#include "guda_b200.h"
*/
import "C"

import "math"

// Reductions of reduce.go:13-64 on device memory.  Each call blocks until the scalar is on the host, like the
// reference's return values.  ArgMax / ArgMin / Max / Min are bit-identical to the reference (NaN and signed-zero
// behaviour of the AVX2 kernels included); Sum and SumSquares use a pairwise tree (ReduceArchTolerance).

func (d DevicePtr) sliceCheck(n int) {
	if n < 0 || n > d.size/4 {
		panic("runtime error: slice bounds out of range") // d.Float32()[:n] in the reference
	}
}

func (d DevicePtr) reduceF(fn func(C.longlong, *C.float, *C.float) C.int, n int, op string) float32 {
	d.sliceCheck(n)
	var r C.float
	if err := status(fn(C.longlong(n), (*C.float)(d.ptr), &r), op); err != nil {
		panic(err)
	}
	return float32(r)
}

// Sum mirrors reduce.go:13-15.
func (d DevicePtr) Sum(n int) float32 {
	return d.reduceF(func(n C.longlong, x, r *C.float) C.int { return C.guda_ssum(n, x, r, nil) }, n, "Sum")
}

// Max mirrors reduce.go:18-23 (-Inf for n == 0).
func (d DevicePtr) Max(n int) float32 {
	if n == 0 {
		return float32(math.Inf(-1))
	}
	return d.reduceF(func(n C.longlong, x, r *C.float) C.int { return C.guda_smax(n, x, r, nil) }, n, "Max")
}

// Min mirrors reduce.go:26-31 (+Inf for n == 0).
func (d DevicePtr) Min(n int) float32 {
	if n == 0 {
		return float32(math.Inf(1))
	}
	return d.reduceF(func(n C.longlong, x, r *C.float) C.int { return C.guda_smin(n, x, r, nil) }, n, "Min")
}

// SumSquares mirrors reduce.go:51-56.
func (d DevicePtr) SumSquares(n int) float32 {
	if n == 0 {
		return 0
	}
	return d.reduceF(func(n C.longlong, x, r *C.float) C.int { return C.guda_ssumsq(n, x, r, nil) }, n, "SumSquares")
}

// Mean mirrors reduce.go:59-64.
func (d DevicePtr) Mean(n int) float32 {
	if n == 0 {
		return 0
	}
	return d.Sum(n) / float32(n)
}

// ArgMax mirrors reduce.go:34-39: first index of the maximum among the non-NaN elements, 0 when x[0] is NaN, -1 for n == 0.
func (d DevicePtr) ArgMax(n int) int {
	if n == 0 {
		return -1
	}
	d.sliceCheck(n)
	var r C.longlong
	if err := status(C.guda_sargmax(C.longlong(n), (*C.float)(d.ptr), &r, nil), "ArgMax"); err != nil {
		panic(err)
	}
	return int(r)
}

// ArgMin mirrors reduce.go:42-47.
func (d DevicePtr) ArgMin(n int) int {
	if n == 0 {
		return -1
	}
	d.sliceCheck(n)
	var r C.longlong
	if err := status(C.guda_sargmin(C.longlong(n), (*C.float)(d.ptr), &r, nil), "ArgMin"); err != nil {
		panic(err)
	}
	return int(r)
}

// Isamax mirrors compute/level1float32.go:89-130: index of the first element with the largest absolute value.
func Isamax(n int, x DevicePtr, incX int) int {
	if incX < 1 {
		if incX == 0 {
			panic("blas: zero x index increment")
		}
		return -1
	}
	if x.size/4 <= (n-1)*incX {
		panic(shortX)
	}
	var r C.longlong
	if err := status(C.guda_isamax(C.int(n), (*C.float)(x.ptr), C.int(incX), &r, nil), "Isamax"); err != nil {
		panic(err)
	}
	return int(r)
}

// Softmax mirrors reduce.go:180-187 (in place over x[:n]): MaxAVX2, fastExp(x - max), sum, ScalUnitary(1/sum).
func Softmax(x DevicePtr, n int) error {
	if n == 0 {
		return nil
	}
	x.sliceCheck(n)
	return status(C.guda_softmax((*C.float)(x.ptr), C.longlong(n), nil), "Softmax")
}

// LayerNorm mirrors fused.go:270-310 (in place over x[:n]); gamma and beta are applied only when both are non-nil
// device pointers, like the reference.
func LayerNorm(x DevicePtr, gamma, beta DevicePtr, n int, eps float32) error {
	x.sliceCheck(n)
	var g, b *C.float
	if gamma.ptr != nil && beta.ptr != nil {
		g, b = (*C.float)(gamma.ptr), (*C.float)(beta.ptr)
	}
	return status(C.guda_layer_norm((*C.float)(x.ptr), g, b, C.longlong(n), C.float(eps), nil), "LayerNorm")
}

// ReduceSum mirrors reduce.go:105: a sum along one axis of a row-major shape[0] x shape[1] matrix (axis 0 -> out[shape[1]],
// axis 1 -> out[shape[0]]).  Every output is one sequential float32 chain in the reference's order, so the bits match.
func ReduceSum(x DevicePtr, shape []int, axis int, out DevicePtr) error {
	return reduceAxis(x, shape, axis, out, false)
}

// ReduceMax mirrors reduce.go:139 (first element, then `val > max`).
func ReduceMax(x DevicePtr, shape []int, axis int, out DevicePtr) error {
	return reduceAxis(x, shape, axis, out, true)
}

func reduceAxis(x DevicePtr, shape []int, axis int, out DevicePtr, isMax bool) error {
	if len(shape) != 2 {
		return ErrInvalidShape
	}
	m, n := shape[0], shape[1]
	if axis != 0 && axis != 1 {
		return ErrInvalidAxis
	}
	outLen := m
	if axis == 0 {
		outLen = n
	}
	checkLen("Reduce", m*n, x)
	checkLen("Reduce", outLen, out)
	if isMax {
		return status(C.guda_reduce_max_axis((*C.float)(x.ptr), C.int(m), C.int(n), C.int(axis), (*C.float)(out.ptr), nil), "ReduceMax")
	}
	return status(C.guda_reduce_sum_axis((*C.float)(x.ptr), C.int(m), C.int(n), C.int(axis), (*C.float)(out.ptr), nil), "ReduceSum")
}

// TopK mirrors reduce.go:247-291: the k largest elements in descending order with their indices, in the order the reference's
// selection sort produces (ties and NaNs included).  k is clamped to n; k <= 0 returns ErrInvalidParameter.
func TopK(x DevicePtr, n, k int) (values []float32, indices []int, err error) {
	if k > n {
		k = n
	}
	if k <= 0 {
		return nil, nil, ErrInvalidParameter
	}
	checkLen("TopK", n, x)
	values = make([]float32, k)
	idx := make([]C.longlong, k)
	if err = status(C.guda_topk((*C.float)(x.ptr), C.longlong(n), C.int(k), (*C.float)(&values[0]), &idx[0], nil), "TopK"); err != nil {
		return nil, nil, err
	}
	indices = make([]int, k)
	for i := range idx {
		indices[i] = int(idx[i])
	}
	return values, indices, nil
}
