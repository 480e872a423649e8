// dmma_lab.cu -- what the FP64 tensor pipe sustains as a function of resident warps and operand pattern (no memory traffic).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/dmma_lab tools/dmma_lab.cu ; ./tools/_bin/dmma_lab
// MODE 0: 8 accumulator tiles, one A and one B register pair for all DMMAs (the bench's peak kernel).
// MODE 1: the DGEMM kernel's register pattern: acc[8][4], af[8], bf[4] -> 32 DMMAs per k4 step, operands changing every step.
// MODE 2: acc[4][4], af[4], bf[4] (the 16-warp layout).
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
template <int MODE>
__global__ void __launch_bounds__(256, 1) lab(const double* __restrict__ in, double* out, int iters) {
  extern __shared__ double sm[];
  constexpr int MF = MODE == 0 ? 8 : (MODE == 1 ? 8 : 4), NF = MODE == 0 ? 1 : 4;
  double acc[MF][NF][2];
#pragma unroll
  for (int i = 0; i < MF; i++)
#pragma unroll
    for (int j = 0; j < NF; j++) { acc[i][j][0] = i; acc[i][j][1] = -j; }
  double af[MF], bf[NF];
#pragma unroll
  for (int i = 0; i < MF; i++) af[i] = in[threadIdx.x + i * 256];
#pragma unroll
  for (int j = 0; j < NF; j++) bf[j] = in[threadIdx.x + (8 + j) * 256];
  for (int it = 0; it < iters; it++) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < MF; i++) dmma(acc[i][0], af[0], bf[0]);
    } else {
#pragma unroll
      for (int i = 0; i < MF; i++)
#pragma unroll
        for (int j = 0; j < NF; j++) dmma(acc[i][j], af[i], bf[j]);
      // next step's fragments: cheap register rotation so that operands differ between steps (stands in for the LDS results)
      const double t = af[0];
#pragma unroll
      for (int i = 0; i + 1 < MF; i++) af[i] = af[i + 1];
      af[MF - 1] = t;
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < MF; i++)
#pragma unroll
    for (int j = 0; j < NF; j++) s += acc[i][j][0] + acc[i][j][1];
  if (s == 12345.678) out[0] = s + sm[0];
}
template <int MODE> void run(const double* in, double* out, int ctas_per_sm, int threads) {
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int smem = ctas_per_sm == 1 ? 200 << 10 : (ctas_per_sm == 2 ? 100 << 10 : (ctas_per_sm == 4 ? 50 << 10 : 24 << 10));
  cudaFuncSetAttribute(lab<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = MODE == 0 ? 40000 : 10000;
  const int per_iter = MODE == 0 ? 8 : (MODE == 1 ? 32 : 16);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0;
  for (int r = 0; r < 3; r++) {
    cudaEventRecord(e0);
    lab<MODE><<<sms * ctas_per_sm, threads, smem>>>(in, out, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double tf = 512.0 * per_iter * (double)iters * (threads / 32) * sms * ctas_per_sm / (ms * 1e-3) / 1e12;
    if (tf > best) best = tf;
  }
  printf("mode %d  %d CTA/SM x %d threads = %2d warps/SM: %.2f TFLOP/s  (%s)\n", MODE, ctas_per_sm, threads, ctas_per_sm * threads / 32, best,
         cudaGetErrorString(cudaGetLastError()));
}
int main() {
  double *in, *out; cudaMalloc(&in, 16 * 256 * 8); cudaMalloc(&out, 64);
  cudaMemset(in, 0x3f, 16 * 256 * 8);
  for (int c : {1, 2, 4, 8}) run<0>(in, out, c, 256);
  for (int c : {1, 2}) run<1>(in, out, c, 256);
  run<1>(in, out, 1, 128);
  for (int c : {1, 2, 4}) run<2>(in, out, c, 256);
  return 0;
}
