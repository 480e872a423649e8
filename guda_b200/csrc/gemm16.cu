// gemm16.cu -- bf16 / fp16 storage types: conversions, 16-bit GEMM entry points, device-side data generator.
// Replaces bfloat16.go:12-29,104-167 (ToBFloat16, GEMMBFloat16) and float16.go:21-87,201-263 (conversions, GEMMFloat16).
//
// Two routes for the 16-bit GEMMs:
//   tensor (default when the shape is eligible): TMA-fed tcgen05 MMA with fp32 TMEM accumulators (gemm_tc.cu);
//           differs from the reference only by fp32 summation order, i.e. by at most one 16-bit ULP after truncation
//   exact : upconvert -> exact-order fp32 SIMT GEMM -> truncate; reproduces the reference's bits
// selected by guda_set_gemm16_mode (GUDA_GEMM16_TENSOR = 0, GUDA_GEMM16_EXACT = 1).
#include "common.cuh"

namespace guda {

int add_bias_act_on(float* x, const float* bias, long long n, int bias_len, cudaStream_t s, bool gelu);
int sgemm_exact_on(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb, float beta,
                   float* c, int ldc, cudaStream_t s);
int sgemm_unfused_noskip(int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb, float beta,
                         float* c, int ldc, cudaStream_t s);
// gemm_tc.cu: returns GUDA_OK / error if handled, -1 if the shape is not eligible for the tensor-core kernel
int gemm16_tc_try(int is_bf16, int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb,
                  float beta, uint16_t* c, int ldc, const float* bias, int epi, cudaStream_t s);

static int g_gemm16_mode = 0;

// ---- bfloat16.go:12-23 ToBFloat16: truncation, +1 only when the low 15 bits are all ones and bit 16 is set ----
__device__ __forceinline__ uint16_t to_bf16_ref(float f) {
  uint32_t bits = __float_as_uint(f);
  uint32_t rounding = (bits >> 16) & 1u;
  if ((bits & 0x7FFFu) == 0x7FFFu && rounding == 1u) bits += 0x10000u;
  return (uint16_t)(bits >> 16);
}
__device__ __forceinline__ float bf16_to_f32_ref(uint16_t b) { return __uint_as_float((uint32_t)b << 16); }

// ---- float16.go:53-87 FromFloat32: truncation; exponent underflow -> signed zero; overflow -> Inf ----
__device__ __forceinline__ uint16_t f32_to_f16_ref(float f) {
  uint32_t bits = __float_as_uint(f);
  uint32_t sign = (bits >> 16) & 0x8000u, exponent = (bits >> 23) & 0xFFu, mant = bits & 0x7FFFFFu;
  if (exponent == 0xFFu) return (uint16_t)(mant == 0 ? (sign | 0x7C00u) : (sign | 0x7C00u | (mant >> 13)));
  int e = (int)exponent - 127 + 15;
  if (e <= 0) return (uint16_t)sign;
  if (e >= 0x1F) return (uint16_t)(sign | 0x7C00u);
  return (uint16_t)(sign | ((uint32_t)e << 10) | (mant >> 13));
}
// ---- float16.go:21-50 ToFloat32, including its subnormal normalisation against bit 9 (restated as written) ----
__device__ __forceinline__ float f16_to_f32_ref(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exponent = (h & 0x7C00u) >> 10, mant = h & 0x3FFu;
  if (exponent == 0) {
    if (mant == 0) return __uint_as_float(sign);
    uint32_t e = 1;
    while ((mant & 0x200u) == 0) { mant <<= 1; e++; }
    mant &= 0x1FFu;
    uint32_t ebits = (uint16_t)(127 - 15 - e + 1);
    return __uint_as_float(sign | (ebits << 23) | (mant << 13));
  }
  if (exponent == 0x1F) return __uint_as_float(mant == 0 ? (sign | 0x7F800000u) : (sign | 0x7FC00000u | (mant << 13)));
  return __uint_as_float(sign | ((exponent + 127 - 15) << 23) | (mant << 13));
}

enum { CVT_F32_BF16 = 0, CVT_BF16_F32 = 1, CVT_F32_F16 = 2, CVT_F16_F32 = 3 };

// strided 2-D convert (rows x cols, separate leading dimensions) so padded matrices convert in one launch
template <int KIND, typename Src, typename Dst>
__global__ void convert_kernel(const Src* __restrict__ src, long long lds, Dst* __restrict__ dst, long long ldd, long long rows, long long cols) {
  long long total = rows * cols;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long r = i / cols, c = i % cols;
    Src v = src[r * lds + c];
    Dst o;
    if (KIND == CVT_F32_BF16) o = to_bf16_ref(v);
    else if (KIND == CVT_BF16_F32) o = bf16_to_f32_ref(v);
    else if (KIND == CVT_F32_F16) o = f32_to_f16_ref(v);
    else o = f16_to_f32_ref(v);
    dst[r * ldd + c] = o;
  }
}

template <int KIND, typename Src, typename Dst>
static int convert2d(const Src* src, long long lds, Dst* dst, long long ldd, long long rows, long long cols, cudaStream_t s) {
  if (rows <= 0 || cols <= 0) return GUDA_OK;
  long long total = rows * cols;
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)sm_count() * 32) blocks = (long long)sm_count() * 32;
  convert_kernel<KIND, Src, Dst><<<(unsigned)blocks, 256, 0, s>>>(src, lds, dst, ldd, rows, cols);
  return check_launch("convert");
}

// ---- counter-based generator: element (r, c) of a logical matrix with `full_cols` columns depends only on
// (seed, r*full_cols + c), so any shard regenerates its block and the host can recompute any sample ----
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
template <typename T>
__global__ void fill_uniform_kernel(T* __restrict__ dst, long long rows, long long cols, long long ld, uint64_t seed, uint64_t row0,
                                    uint64_t col0, uint64_t full_cols, T lo, T hi) {
  long long total = rows * cols;
  T scale = hi - lo;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long r = i / cols, c = i % cols;
    uint64_t idx = (row0 + (uint64_t)r) * full_cols + col0 + (uint64_t)c;
    uint64_t h = mix64(seed * 0xD1342543DE82EF95ull + idx);
    T u;
    if (sizeof(T) == 4) u = (T)((float)(h >> 40) * (1.0f / 16777216.0f));
    else u = (T)((double)(h >> 11) * (1.0 / 9007199254740992.0));
    T v;
    if (sizeof(T) == 4) v = (T)__fadd_rn(__fmul_rn((float)u, (float)scale), (float)lo);
    else v = (T)__dadd_rn(__dmul_rn((double)u, (double)scale), (double)lo);
    dst[r * ld + c] = v;
  }
}


static int gemm16_impl(int is_bf16, int ta, int tb, int m, int n, int k, float alpha, const uint16_t* a, int lda,
                       const uint16_t* b, int ldb, float beta, uint16_t* c, int ldc, const float* bias, int epi,
                       guda_stream_t stream, const char* op) {
  if (m < 0) GUDA_PANIC("blas: m < 0");
  if (n < 0) GUDA_PANIC("blas: n < 0");
  if (k < 0) GUDA_PANIC("blas: k < 0");
  if (is_bf16 && (ta || tb))   // bfloat16.go:104 silently ignores the flags; refusing is safer than a silent wrong answer
    return set_error(GUDA_ERR_NOT_IMPLEMENTED, "%s: transposed operands are not supported (the reference ignores the flags)", op);
  auto mx = [](int x, int y) { return x > y ? x : y; };
  if (lda < mx(1, ta ? m : k)) GUDA_PANIC("blas: bad leading dimension of A");
  if (ldb < mx(1, tb ? k : n)) GUDA_PANIC("blas: bad leading dimension of B");
  if (ldc < mx(1, n)) GUDA_PANIC("blas: bad leading dimension of C");
  if (m == 0 || n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!c || ((!a || !b) && k > 0)) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", op);
  cudaStream_t s = native(stream);

  if (g_gemm16_mode == 0 && !ta && !tb) {
    rc = gemm16_tc_try(is_bf16, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, epi, s);
    if (rc >= 0) return rc;
  }
  // exact route: fp32 temporaries from the pool (stream-ordered use; freed blocks stay cached in the pool)
  const int ar = ta ? k : m, ac = ta ? m : k, br = tb ? n : k, bc = tb ? k : n;
  PoolTemp af, bf, cf;
  if ((rc = af.alloc((size_t)ar * ac * 4 + 16, s))) return rc;
  if ((rc = bf.alloc((size_t)br * bc * 4 + 16, s))) return rc;
  if ((rc = cf.alloc((size_t)m * n * 4 + 16, s))) return rc;
  float *A = (float*)af.p, *B = (float*)bf.p, *C = (float*)cf.p;
  if (is_bf16) {
    if ((rc = convert2d<CVT_BF16_F32>(a, lda, A, ac, ar, ac, s))) return rc;
    if ((rc = convert2d<CVT_BF16_F32>(b, ldb, B, bc, br, bc, s))) return rc;
    if (beta != 0 && (rc = convert2d<CVT_BF16_F32>(c, ldc, C, n, m, n, s))) return rc;
    if ((rc = sgemm_unfused_noskip(m, n, k, alpha, A, ac, B, bc, beta, C, n, s))) return rc;
    if (epi == 1 && (rc = add_bias_act_on(C, bias, (long long)m * n, n, s, false))) return rc;
    if (epi == 2 && (rc = add_bias_act_on(C, bias, (long long)m * n, n, s, true))) return rc;
    rc = convert2d<CVT_F32_BF16>(C, n, c, ldc, m, n, s);
  } else {
    if ((rc = convert2d<CVT_F16_F32>(a, lda, A, ac, ar, ac, s))) return rc;
    if ((rc = convert2d<CVT_F16_F32>(b, ldb, B, bc, br, bc, s))) return rc;
    if (beta != 0) { if ((rc = convert2d<CVT_F16_F32>(c, ldc, C, n, m, n, s))) return rc; }
    else GUDA_CUDA(cudaMemsetAsync(C, 0, (size_t)m * n * 4, s));
    if ((rc = sgemm_exact_on(ta, tb, m, n, k, alpha, A, ac, B, bc, beta, C, n, s))) return rc;      // float16.go:231 (exact order whatever the SGEMM math mode)
    rc = convert2d<CVT_F32_F16>(C, n, c, ldc, m, n, s);
  }
  if (rc) return rc;
  // the temporaries go back to the pool here; the pool orders their next owner after this stream's queued work
  return GUDA_OK;
}

}  // namespace guda

using namespace guda;
extern "C" {

int guda_set_gemm16_mode(int mode) {
  if (mode != 0 && mode != 1) return set_error(GUDA_ERR_INVALID_ARG, "unknown gemm16 mode %d", mode);
  g_gemm16_mode = mode;
  return GUDA_OK;
}
int guda_get_gemm16_mode(int* mode) { *mode = g_gemm16_mode; return GUDA_OK; }

int guda_gemm_bf16(int ta, int tb, int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb,
                   float beta, uint16_t* c, int ldc, guda_stream_t stream) {
  return gemm16_impl(1, ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, nullptr, 0, stream, "GEMMBFloat16");
}
int guda_gemm_f16(int ta, int tb, int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb,
                  float beta, uint16_t* c, int ldc, guda_stream_t stream) {
  return gemm16_impl(0, ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, nullptr, 0, stream, "GEMMFloat16");
}
int guda_gemm_bf16_bias_relu(int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb,
                             uint16_t* c, int ldc, const float* bias, guda_stream_t stream) {
  if (!bias) return set_error(GUDA_ERR_INVALID_ARG, "null bias");
  return gemm16_impl(1, 0, 0, m, n, k, alpha, a, lda, b, ldb, 0.0f, c, ldc, bias, 1, stream, "GEMMBFloat16BiasReLU");
}
int guda_gemm_bf16_bias_gelu(int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb,
                             uint16_t* c, int ldc, const float* bias, guda_stream_t stream) {
  if (!bias) return set_error(GUDA_ERR_INVALID_ARG, "null bias");
  return gemm16_impl(1, 0, 0, m, n, k, alpha, a, lda, b, ldb, 0.0f, c, ldc, bias, 2, stream, "GEMMBFloat16BiasGELU");
}

#define GUDA_CVT(NAME, KIND, SRC, DST)                                                    \
  int NAME(const SRC* src, DST* dst, size_t n, guda_stream_t stream) {                    \
    int rc = bind_device();                                                               \
    if (rc) return rc;                                                                    \
    if (n == 0) return GUDA_OK;                                                           \
    if (!src || !dst) return set_error(GUDA_ERR_INVALID_ARG, #NAME ": null pointer");     \
    return convert2d<KIND>(src, (long long)n, dst, (long long)n, 1, (long long)n, native(stream)); \
  }
GUDA_CVT(guda_f32_to_bf16, CVT_F32_BF16, float, uint16_t)
GUDA_CVT(guda_bf16_to_f32, CVT_BF16_F32, uint16_t, float)
GUDA_CVT(guda_f32_to_f16, CVT_F32_F16, float, uint16_t)
GUDA_CVT(guda_f16_to_f32, CVT_F16_F32, uint16_t, float)

int guda_fill_uniform_f32(float* dst, size_t rows, size_t cols, size_t ld, uint64_t seed, uint64_t row0, uint64_t col0,
                          uint64_t full_cols, float lo, float hi, guda_stream_t stream) {
  int rc = bind_device();
  if (rc) return rc;
  if (rows == 0 || cols == 0) return GUDA_OK;
  if (!dst || ld < cols) return set_error(GUDA_ERR_INVALID_ARG, "fill: bad arguments");
  long long blocks = ((long long)(rows * cols) + 255) / 256;
  if (blocks > (long long)sm_count() * 32) blocks = (long long)sm_count() * 32;
  fill_uniform_kernel<float><<<(unsigned)blocks, 256, 0, native(stream)>>>(dst, rows, cols, ld, seed, row0, col0, full_cols, lo, hi);
  return check_launch("fill");
}
int guda_fill_uniform_f64(double* dst, size_t rows, size_t cols, size_t ld, uint64_t seed, uint64_t row0, uint64_t col0,
                          uint64_t full_cols, double lo, double hi, guda_stream_t stream) {
  int rc = bind_device();
  if (rc) return rc;
  if (rows == 0 || cols == 0) return GUDA_OK;
  if (!dst || ld < cols) return set_error(GUDA_ERR_INVALID_ARG, "fill: bad arguments");
  long long blocks = ((long long)(rows * cols) + 255) / 256;
  if (blocks > (long long)sm_count() * 32) blocks = (long long)sm_count() * 32;
  fill_uniform_kernel<double><<<(unsigned)blocks, 256, 0, native(stream)>>>(dst, rows, cols, ld, seed, row0, col0, full_cols, lo, hi);
  return check_launch("fill");
}

}  // extern "C"
