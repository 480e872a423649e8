// microbench.cu -- measured denominators the driver's MEASURED_PEAKS.json does not carry:
// FP32 FFMA and FP64 DFMA SIMT peaks (the pipes the exact-order GEMMs are bound by) and a
// 128-bit streaming copy (the Level-1 bound).  Used by bench.py live, next to each roofline.
#include "common.cuh"

namespace guda {

template <typename T>
__global__ void __launch_bounds__(256) fma_peak_kernel(T* out, int iters, T seed) {
  T a0 = seed + threadIdx.x, a1 = a0 + T(1), a2 = a0 + T(2), a3 = a0 + T(3), a4 = a0 + T(4), a5 = a0 + T(5), a6 = a0 + T(6), a7 = a0 + T(7);
  const T x = T(0.999), y = T(1e-3);
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      a0 = fma(a0, x, y); a1 = fma(a1, x, y); a2 = fma(a2, x, y); a3 = fma(a3, x, y);
      a4 = fma(a4, x, y); a5 = fma(a5, x, y); a6 = fma(a6, x, y); a7 = fma(a7, x, y);
    }
  }
  T s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (s == T(12345.678)) out[0] = s;       // never true; keeps the chain alive
}

// packed fp32: fma.rn.f32x2 (FFMA2), 8 independent 2-wide chains per thread
__global__ void __launch_bounds__(256) ffma2_peak_kernel(float* out, int iters) {
  unsigned long long a[8], x, y;
  const float xf = 0.999f, yf = 1e-3f;
  asm("mov.b64 %0, {%1, %1};" : "=l"(x) : "f"(xf));
  asm("mov.b64 %0, {%1, %1};" : "=l"(y) : "f"(yf));
#pragma unroll
  for (int i = 0; i < 8; i++) { float f = 1.0f + threadIdx.x + i; asm("mov.b64 %0, {%1, %1};" : "=l"(a[i]) : "f"(f)); }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 8; u++)
#pragma unroll
      for (int i = 0; i < 8; i++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[i]) : "l"(x), "l"(y));
  }
  unsigned long long r = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) r ^= a[i];
  if (r == 0x12345678ull) out[0] = 1.0f;
}

// FP64 tensor pipe: mma.sync.m8n8k4.f64 (DMMA), 8 independent accumulator chains per warp
__global__ void __launch_bounds__(256) dmma_peak_kernel(double* out, int iters) {
  double a = 1.0 + threadIdx.x * 1e-3, b = 0.999;
  double c[8][2];
#pragma unroll
  for (int i = 0; i < 8; i++) { c[i][0] = i; c[i][1] = -i; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1];
  if (s == 12345.678) out[0] = s;
}

__global__ void __launch_bounds__(256) copy_peak_kernel(const float4* __restrict__ src, float4* __restrict__ dst, size_t nvec) {
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride * 4) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) if (i + u * stride < nvec) v[u] = __ldcs(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; u++) if (i + u * stride < nvec) __stcs(dst + i + u * stride, v[u]);
  }
}

template <typename T>
static int fma_peak(double* tflops, int reps) {
  int rc = bind_device();
  if (rc) return rc;
  T* out;
  GUDA_CUDA(cudaMalloc((void**)&out, 256));
  cudaEvent_t e0, e1;
  GUDA_CUDA(cudaEventCreate(&e0)); GUDA_CUDA(cudaEventCreate(&e1));
  const int iters = sizeof(T) == 4 ? 4096 : 2048;
  const int blocks = sm_count() * 8;
  cudaStream_t s = native(nullptr);
  double best = 0;
  for (int r = 0; r < reps + 1; r++) {
    GUDA_CUDA(cudaEventRecord(e0, s));
    fma_peak_kernel<T><<<blocks, 256, 0, s>>>(out, iters, T(1));
    GUDA_CUDA(cudaEventRecord(e1, s));
    GUDA_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    GUDA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    double flops = 2.0 * 64.0 * iters * 256.0 * blocks;
    double t = flops / (ms * 1e-3) / 1e12;
    if (r > 0 && t > best) best = t;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
  *tflops = best;
  return GUDA_OK;
}

}  // namespace guda

using namespace guda;
extern "C" {
// kind: 0 = FP32 FFMA TFLOP/s, 1 = FP64 DFMA TFLOP/s, 2 = 128-bit copy GB/s (read+write bytes, 1 GiB each way)
int guda_microbench_peak(int kind, int reps, double* value) {
  if (!value) return set_error(GUDA_ERR_INVALID_ARG, "null pointer");
  if (reps < 1) reps = 3;
  if (kind == 0) return fma_peak<float>(value, reps);
  if (kind == 1) return fma_peak<double>(value, reps);
  int rc = bind_device();
  if (rc) return rc;
  if (kind == 4) {            // packed fp32 FMA (FFMA2) TFLOP/s
    float* out;
    GUDA_CUDA(cudaMalloc((void**)&out, 256));
    cudaEvent_t e0, e1;
    GUDA_CUDA(cudaEventCreate(&e0)); GUDA_CUDA(cudaEventCreate(&e1));
    const int iters = 4096, blocks = sm_count() * 8;
    cudaStream_t s = native(nullptr);
    double best = 0;
    for (int r = 0; r < reps + 1; r++) {
      GUDA_CUDA(cudaEventRecord(e0, s));
      ffma2_peak_kernel<<<blocks, 256, 0, s>>>(out, iters);
      GUDA_CUDA(cudaEventRecord(e1, s));
      GUDA_CUDA(cudaEventSynchronize(e1));
      float ms = 0;
      GUDA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      double t = 2.0 * 2.0 * 64.0 * iters * 256.0 * blocks / (ms * 1e-3) / 1e12;
      if (r > 0 && t > best) best = t;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    *value = best;
    return GUDA_OK;
  }
  if (kind == 3) {            // FP64 tensor pipe (DMMA m8n8k4) TFLOP/s
    double* out;
    GUDA_CUDA(cudaMalloc((void**)&out, 256));
    cudaEvent_t e0, e1;
    GUDA_CUDA(cudaEventCreate(&e0)); GUDA_CUDA(cudaEventCreate(&e1));
    const int iters = 2048, blocks = sm_count() * 8;
    cudaStream_t s = native(nullptr);
    double best = 0;
    for (int r = 0; r < reps + 1; r++) {
      GUDA_CUDA(cudaEventRecord(e0, s));
      dmma_peak_kernel<<<blocks, 256, 0, s>>>(out, iters);
      GUDA_CUDA(cudaEventRecord(e1, s));
      GUDA_CUDA(cudaEventSynchronize(e1));
      float ms = 0;
      GUDA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      double t = 512.0 * 8.0 * iters * 8.0 * blocks / (ms * 1e-3) / 1e12;
      if (r > 0 && t > best) best = t;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    *value = best;
    return GUDA_OK;
  }
  if (kind != 2) return set_error(GUDA_ERR_INVALID_ARG, "unknown microbenchmark %d", kind);
  const size_t bytes = (size_t)1 << 30;
  float4 *src, *dst;
  GUDA_CUDA(cudaMalloc((void**)&src, bytes));
  GUDA_CUDA(cudaMalloc((void**)&dst, bytes));
  GUDA_CUDA(cudaMemset(src, 1, bytes));
  cudaEvent_t e0, e1;
  GUDA_CUDA(cudaEventCreate(&e0)); GUDA_CUDA(cudaEventCreate(&e1));
  cudaStream_t s = native(nullptr);
  double best = 0;
  for (int r = 0; r < reps + 1; r++) {
    GUDA_CUDA(cudaEventRecord(e0, s));
    copy_peak_kernel<<<sm_count() * 16, 256, 0, s>>>(src, dst, bytes / 16);
    GUDA_CUDA(cudaEventRecord(e1, s));
    GUDA_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    GUDA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    double gbs = 2.0 * bytes / (ms * 1e-3) / 1e9;
    if (r > 0 && gbs > best) best = gbs;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(src); cudaFree(dst);
  *value = best;
  return GUDA_OK;
}

}
