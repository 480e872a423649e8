package guda

/*
#include "guda_b200.h"
*/
import "C"

import "unsafe"

// Add, Multiply, ReLU and Sigmoid mirror math.go:133-169 and :230-272 on device memory: one streaming pass each, every
// element computed with the reference's float32 statements (Sigmoid is activations.go:9-27 SigmoidFloat32).
func Add(a, b, c DevicePtr, n int) error {
	checkLen("Add", n, a, b, c)
	return status(C.guda_add((*C.float)(a.ptr), (*C.float)(b.ptr), (*C.float)(c.ptr), C.longlong(n), nil), "Add")
}

func Multiply(a, b, c DevicePtr, n int) error {
	checkLen("Multiply", n, a, b, c)
	return status(C.guda_multiply((*C.float)(a.ptr), (*C.float)(b.ptr), (*C.float)(c.ptr), C.longlong(n), nil), "Multiply")
}

func ReLU(x DevicePtr, n int) error {
	checkLen("ReLU", n, x)
	return status(C.guda_relu((*C.float)(x.ptr), C.longlong(n), nil), "ReLU")
}

func Sigmoid(x DevicePtr, n int) error {
	checkLen("Sigmoid", n, x)
	return status(C.guda_sigmoid((*C.float)(x.ptr), C.longlong(n), nil), "Sigmoid")
}

// checkLen panics like the reference's slice indexing would when n runs past a DevicePtr.
func checkLen(op string, n int, ptrs ...DevicePtr) {
	for _, p := range ptrs {
		if n < 0 || p.size/4 < n {
			panic("runtime error: index out of range in " + op)
		}
	}
}

// FusedOpType and FusedKernel mirror fused.go:18-118.  Custom accepts the one function the reference itself passes to it
// (geluFloat32, fused.go:261-267): a Go closure cannot run on the device, so any other function returns ErrNotSupported
// from Execute.
type FusedOpType int

const (
	FusedAdd FusedOpType = iota
	FusedMul
	FusedAddScalar
	FusedMulScalar
	FusedReLU
	FusedSigmoid
	FusedCustom
)

type FusedOp struct {
	Type         FusedOpType
	Alpha        float32
	Broadcast    bool
	BroadcastDim int
	gelu         bool
}

type FusedKernel struct {
	ops []FusedOp
}

func NewFusedKernel() *FusedKernel { return &FusedKernel{ops: make([]FusedOp, 0, 4)} }

func (fk *FusedKernel) Add(alpha float32) *FusedKernel {
	fk.ops = append(fk.ops, FusedOp{Type: FusedAdd, Alpha: alpha, BroadcastDim: -1})
	return fk
}
func (fk *FusedKernel) AddBroadcast(alpha float32, broadcastDim int) *FusedKernel {
	fk.ops = append(fk.ops, FusedOp{Type: FusedAdd, Alpha: alpha, Broadcast: true, BroadcastDim: broadcastDim})
	return fk
}
func (fk *FusedKernel) Multiply() *FusedKernel {
	fk.ops = append(fk.ops, FusedOp{Type: FusedMul, BroadcastDim: -1})
	return fk
}
func (fk *FusedKernel) AddScalar(alpha float32) *FusedKernel {
	fk.ops = append(fk.ops, FusedOp{Type: FusedAddScalar, Alpha: alpha})
	return fk
}
func (fk *FusedKernel) MulScalar(alpha float32) *FusedKernel {
	fk.ops = append(fk.ops, FusedOp{Type: FusedMulScalar, Alpha: alpha})
	return fk
}
func (fk *FusedKernel) ReLU() *FusedKernel {
	fk.ops = append(fk.ops, FusedOp{Type: FusedReLU})
	return fk
}
func (fk *FusedKernel) Sigmoid() *FusedKernel {
	fk.ops = append(fk.ops, FusedOp{Type: FusedSigmoid})
	return fk
}

// GeluFloat32 stands for fused_gelu.go:44 geluFloat32 as an argument of Custom; it is evaluated on the device.
var GeluFloat32 = func(float32) float32 { panic("guda: geluFloat32 runs on the device; pass it to FusedKernel.Custom") }

func (fk *FusedKernel) Custom(fn func(float32) float32) *FusedKernel {
	isGelu := fn != nil && *(*unsafe.Pointer)(unsafe.Pointer(&fn)) == *(*unsafe.Pointer)(unsafe.Pointer(&GeluFloat32))
	fk.ops = append(fk.ops, FusedOp{Type: FusedCustom, gelu: isGelu})
	return fk
}

// Execute mirrors fused.go:123-216: one read of x[idx], the ops in order, one write.
func (fk *FusedKernel) Execute(x DevicePtr, others []DevicePtr, n int, shapes ...[]int) error {
	checkLen("FusedOp", n, x)
	if len(fk.ops) > C.GUDA_FUSED_MAX_OPS || len(others) > C.GUDA_FUSED_MAX_OTHERS {
		return ErrInvalidParameter
	}
	var ops [C.GUDA_FUSED_MAX_OPS]C.guda_fused_op_t
	var ptrs [C.GUDA_FUSED_MAX_OTHERS]*C.float
	for i, o := range others {
		ptrs[i] = (*C.float)(o.ptr)
	}
	oi := 0
	for k, op := range fk.ops {
		ops[k]._type = C.int(op.Type)
		ops[k].alpha = C.float(op.Alpha)
		if op.Type == FusedCustom {
			if !op.gelu {
				return ErrNotSupported
			}
			ops[k]._type = C.GUDA_FUSED_GELU
		}
		if (op.Type == FusedAdd || op.Type == FusedMul) && oi < len(others) {
			if op.Broadcast && op.BroadcastDim >= 0 && oi < len(shapes) {
				shape := shapes[oi]
				if op.BroadcastDim >= len(shape) { // fused.go:155: the operand silently becomes 0 (Add) / 1 (Mul)
					return ErrNotSupported
				}
				size := 1
				for _, d := range shape[op.BroadcastDim:] {
					size *= d
				}
				dim := shape[op.BroadcastDim]
				if dim <= 0 || others[oi].size/4 < dim {
					return ErrInvalidShape
				}
				ops[k].broadcast_size = C.longlong(size)
				ops[k].broadcast_dim = C.longlong(dim)
			} else {
				checkLen("FusedOp", n, others[oi])
			}
			oi++
		}
	}
	var pp **C.float
	if len(others) > 0 {
		pp = &ptrs[0]
	}
	return status(C.guda_fused_execute((*C.float)(x.ptr), pp, C.int(len(others)), C.longlong(n), &ops[0], C.int(len(fk.ops)), nil), "FusedOp")
}

// LinearReLU mirrors fused.go:252: x = ReLU(alpha*x + beta).
func LinearReLU(x DevicePtr, alpha, beta float32, n int) error {
	return NewFusedKernel().MulScalar(alpha).AddScalar(beta).ReLU().Execute(x, nil, n)
}

// GELU mirrors fused.go:261: x = geluFloat32(x) in place.
func GELU(x DevicePtr, n int) error {
	return NewFusedKernel().Custom(GeluFloat32).Execute(x, nil, n)
}
