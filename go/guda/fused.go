package guda

/*
#include "guda_b200.h"
*/
import "C"

// FusedGEMMBiasReLU mirrors fused.go:313-341: C = max(0, alpha*A*B + bias[j]); transposes or beta != 0
// return ErrNotSupported.  One kernel: bias and activation are applied to the accumulators in registers
// before the single store of C (the reference makes a second, column-order pass over C).
func FusedGEMMBiasReLU(transA, transB bool, m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int,
	beta float32, c DevicePtr, ldc int, bias DevicePtr) error {
	return status(C.guda_sgemm_bias_relu(cbool(transA), cbool(transB), C.int(m), C.int(n), C.int(k), C.float(alpha),
		(*C.float)(a.ptr), C.int(lda), (*C.float)(b.ptr), C.int(ldb), C.float(beta), (*C.float)(c.ptr), C.int(ldc),
		(*C.float)(bias.ptr), nil), "FusedGEMMBiasReLU")
}

// FusedGEMMBiasGELU mirrors fused_gelu.go:65-93 with the scalar route's geluFloat32 (fused_gelu.go:44-56) bit for bit.
func FusedGEMMBiasGELU(transA, transB bool, m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int,
	beta float32, c DevicePtr, ldc int, bias DevicePtr) error {
	return status(C.guda_sgemm_bias_gelu(cbool(transA), cbool(transB), C.int(m), C.int(n), C.int(k), C.float(alpha),
		(*C.float)(a.ptr), C.int(lda), (*C.float)(b.ptr), C.int(ldb), C.float(beta), (*C.float)(c.ptr), C.int(ldc),
		(*C.float)(bias.ptr), nil), "FusedGEMMBiasGELU")
}

// AddBiasGELU / AddBiasReLU mirror fused_gelu.go:12-41 and fused.go:220-260: x[i] = act(x[i] + bias[i % len(bias)]).
func AddBiasGELU(x, bias DevicePtr, n, biasLen int) error {
	return status(C.guda_add_bias_gelu((*C.float)(x.ptr), (*C.float)(bias.ptr), C.int(n), C.int(biasLen), nil), "AddBiasGELU")
}
func AddBiasReLU(x, bias DevicePtr, n, biasLen int) error {
	return status(C.guda_add_bias_relu((*C.float)(x.ptr), (*C.float)(bias.ptr), C.int(n), C.int(biasLen), nil), "AddBiasReLU")
}

// GEMMBFloat16 mirrors bfloat16.go:104-167 (bf16 storage, fp32 accumulation).  The reference ignores the
// transpose flags; here they must be false.  Default route: tcgen05 tensor cores, within 1 bf16 ULP of the
// reference; SetGemm16Mode(1) selects the route that reproduces its bits.
func GEMMBFloat16(transA, transB bool, m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int,
	beta float32, c DevicePtr, ldc int) error {
	return status(C.guda_gemm_bf16(cbool(transA), cbool(transB), C.int(m), C.int(n), C.int(k), C.float(alpha),
		(*C.uint16_t)(a.ptr), C.int(lda), (*C.uint16_t)(b.ptr), C.int(ldb), C.float(beta), (*C.uint16_t)(c.ptr), C.int(ldc), nil), "GEMMBFloat16")
}

// GEMMFloat16 mirrors float16.go:201-244.
func GEMMFloat16(transA, transB bool, m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int,
	beta float32, c DevicePtr, ldc int) error {
	return status(C.guda_gemm_f16(cbool(transA), cbool(transB), C.int(m), C.int(n), C.int(k), C.float(alpha),
		(*C.uint16_t)(a.ptr), C.int(lda), (*C.uint16_t)(b.ptr), C.int(ldb), C.float(beta), (*C.uint16_t)(c.ptr), C.int(ldc), nil), "GEMMFloat16")
}

// FusedGEMMBFloat16BiasGELU / ReLU are new: the C3 FFN shape in bf16 with the epilogue fused into the tcgen05 kernel.
func FusedGEMMBFloat16BiasGELU(m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int, c DevicePtr, ldc int, bias DevicePtr) error {
	return status(C.guda_gemm_bf16_bias_gelu(C.int(m), C.int(n), C.int(k), C.float(alpha), (*C.uint16_t)(a.ptr), C.int(lda),
		(*C.uint16_t)(b.ptr), C.int(ldb), (*C.uint16_t)(c.ptr), C.int(ldc), (*C.float)(bias.ptr), nil), "FusedGEMMBFloat16BiasGELU")
}
func FusedGEMMBFloat16BiasReLU(m, n, k int, alpha float32, a DevicePtr, lda int, b DevicePtr, ldb int, c DevicePtr, ldc int, bias DevicePtr) error {
	return status(C.guda_gemm_bf16_bias_relu(C.int(m), C.int(n), C.int(k), C.float(alpha), (*C.uint16_t)(a.ptr), C.int(lda),
		(*C.uint16_t)(b.ptr), C.int(ldb), (*C.uint16_t)(c.ptr), C.int(ldc), (*C.float)(bias.ptr), nil), "FusedGEMMBFloat16BiasReLU")
}

func SetGemm16Mode(mode int) error { return status(C.guda_set_gemm16_mode(C.int(mode)), "SetGemm16Mode") }
func SetDgemmMode(mode int) error  { return status(C.guda_set_dgemm_mode(C.int(mode)), "SetDgemmMode") }
