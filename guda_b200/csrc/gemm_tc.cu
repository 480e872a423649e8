// gemm_tc.cu -- tcgen05 / TMEM / TMA tensor-core GEMM paths (bf16, fp16, 3xTF32).
#include "common.cuh"
namespace guda {
int sgemm_3xtf32_try(int, int, int, int, int, float, const float*, int, const float*, int, float, float*, int, cudaStream_t) { return -1; }
int gemm16_tc_try(int, int, int, int, float, const uint16_t*, int, const uint16_t*, int, float, uint16_t*, int, const float*, int,
                  cudaStream_t) { return -1; }
}  // namespace guda
