/* fused.c -- oracle (TEST INFRASTRUCTURE): activations, fused GEMM epilogues,
 * bf16/fp16 storage types, generators and the tolerance framework.  See guda_oracle.h. */
#include "guda_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* ---------------------------------------------------------------- activations */

/* activations.go:60-90 ExpFloat32.  Go evaluates float32 expressions without
 * fusing on amd64; untyped constants convert to float32 where they meet a float32. */
float orc_exp_f32(float x) {
  if (x > 88.7f) return 3.40282346638528859811704183484516925440e+38f;   /* math.MaxFloat32 */
  if (x < -87.3f) return 0.0f;
  const double ln2 = 0.6931471805599453094;                              /* math_constants.go:20 */
  int k = (int)floor((double)x / ln2);
  float kf = (float)k * (float)ln2;
  float r = x - kf;
  float r2 = r * r;
  float r3 = r2 * r;
  float r4 = r2 * r2;
  float r5 = r4 * r;
  float e = 1.0f + r;
  float t;
  t = 0.4999999701976776f * r2; e = e + t;
  t = 0.1666666567325592f * r3; e = e + t;
  t = 0.0416666679084301f * r4; e = e + t;
  t = 0.0083333337679505f * r5; e = e + t;
  return (float)ldexp((double)e, k);
}

/* activations.go:31-56 TanhFloat32 (NaN would recurse forever in the reference; we return NaN) */
float orc_tanh_f32(float x) {
  if (x != x) return x;
  if (x > 10.0f) return 1.0f;
  if (x < -10.0f) return -1.0f;
  if (x >= 0) {
    if (x < 0.5f) {
      float x2 = x * x;
      float a = x2 / 3.0f;
      float b = 1.0f - a;
      float c = 2.0f * x2; c = c * x2; c = c / 15.0f;
      float s = b + c;
      return x * s;
    }
    float e = orc_exp_f32(2.0f * x);
    float num = e - 1.0f, den = e + 1.0f;
    return num / den;
  }
  return -orc_tanh_f32(-x);
}

/* fused_gelu.go:44-56 geluFloat32; constants math_constants.go:31-33 */
float orc_gelu_f32(float x) {
  float x3 = x * x; x3 = x3 * x;
  float inner = 0.044715f * x3; inner = x + inner;
  float arg = 0.7978845608028653559f * inner;
  float t = orc_tanh_f32(arg);
  float h = 0.5f * x;
  float u = 1.0f + t;
  return h * u;
}

/* compute/asm/f32/gelu_avx2_amd64.s:46-77 (vector body; the scalar tail :100-135 is the same maths) */
float orc_gelu_pade(float x) {
  const float c_s2pi = u2f(0x3f4c422a), c_coef = u2f(0x3d372713);
  float x2 = x * x;
  float x3 = x2 * x;
  float t = c_coef * x3;
  t = t + x;
  float arg = c_s2pi * t;
  arg = (arg > -3.0f) ? arg : -3.0f;       /* VMAXPS: NaN -> second source (-3) */
  arg = (arg < 3.0f) ? arg : 3.0f;
  float a2 = arg * arg;
  float n9 = 9.0f * a2;
  float d1 = 27.0f + a2;
  float d2 = 27.0f + n9;
  float num = arg * d1;
  float th = num / d2;
  th = th + 1.0f;
  float h = 0.5f * x;
  return h * th;
}

/* reference.go:203-211 Reference.GELU (float64 tanh) */
void orc_ref_gelu(float* x, size_t n) {
  for (size_t i = 0; i < n; i++) {
    float x3 = x[i] * x[i]; x3 = x3 * x[i];
    double arg = sqrt(2.0 / 3.14159265358979323846) * ((double)x[i] + 0.044715 * (double)x3);
    float th = (float)tanh(arg);
    float h = 0.5f * x[i];
    float u = 1.0f + th;
    x[i] = h * u;
  }
}

/* fused_gelu.go:12-41 AddBiasGELU: x[i] = gelu(x[i] + bias[i % bias_len]) */
void orc_add_bias_gelu(float* x, const float* bias, int n, int bias_len) {
  for (int i = 0; i < n; i++) { float v = x[i] + bias[i % bias_len]; x[i] = orc_gelu_f32(v); }
}

/* ---------------------------------------------------------------- fused GEMM */

/* fused.go:313-368: trans or beta != 0 -> ErrNotSupported; stock blas32.Gemm with
 * beta = 0; then a second pass c = max(0, c + bias[j]). */
int orc_fused_gemm_bias_relu(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda,
                             const float* b, int ldb, float beta, float* c, int ldc, const float* bias, int threads) {
  if (ta || tb) return ORC_NOT_SUPPORTED;
  if (beta != 0) return ORC_NOT_SUPPORTED;
  int rc = orc_sgemm(0, 0, m, n, k, alpha, a, (size_t)(m ? (size_t)(m - 1) * lda + k : 0), lda,
                     b, (size_t)(k ? (size_t)(k - 1) * ldb + n : 0), ldb, 0.0f,
                     c, (size_t)(m ? (size_t)(m - 1) * ldc + n : 0), ldc, 1, threads);
  if (rc) return rc;
  for (int j = 0; j < n; j++) {
    float bv = bias[j];
    for (int i = 0; i < m; i++) {
      float v = c[(size_t)i * ldc + j] + bv;
      if (v < 0) v = 0;
      c[(size_t)i * ldc + j] = v;
    }
  }
  return ORC_OK;
}

/* fused_gelu.go:65-137 */
int orc_fused_gemm_bias_gelu(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda,
                             const float* b, int ldb, float beta, float* c, int ldc, const float* bias,
                             int gelu_route, int threads) {
  if (ta || tb) return ORC_NOT_SUPPORTED;
  if (beta != 0) return ORC_NOT_SUPPORTED;
  int rc = orc_sgemm(0, 0, m, n, k, alpha, a, (size_t)(m ? (size_t)(m - 1) * lda + k : 0), lda,
                     b, (size_t)(k ? (size_t)(k - 1) * ldb + n : 0), ldb, 0.0f,
                     c, (size_t)(m ? (size_t)(m - 1) * ldc + n : 0), ldc, 1, threads);
  if (rc) return rc;
  for (int j = 0; j < n; j++) {
    float bv = bias[j];
    for (int i = 0; i < m; i++) {
      float v = c[(size_t)i * ldc + j] + bv;
      c[(size_t)i * ldc + j] = gelu_route ? orc_gelu_pade(v) : orc_gelu_f32(v);
    }
  }
  return ORC_OK;
}

/* ---------------------------------------------------------------- bf16 / fp16 */

/* bfloat16.go:12-23: documented as RNE but the first test is dead code, so this
 * is truncation except when the low 15 bits are all ones and bit 16 is set. */
uint16_t orc_to_bf16(float f) {
  uint32_t bits = f2u(f);
  uint32_t rounding = (bits >> 16) & 1;
  if ((bits & 0x7FFF) > 0x7FFF || ((bits & 0x7FFF) == 0x7FFF && rounding == 1)) bits += 0x10000;
  return (uint16_t)(bits >> 16);
}
float orc_bf16_to_f32(uint16_t b) { return u2f((uint32_t)b << 16); }     /* bfloat16.go:26-29 */

/* float16.go:53-87 FromFloat32: truncation, subnormal results flush to signed zero */
uint16_t orc_f32_to_f16(float f) {
  uint32_t bits = f2u(f);
  uint32_t sign = (bits >> 16) & 0x8000, exponent = (bits >> 23) & 0xFF, mant = bits & 0x7FFFFF;
  if (exponent == 0xFF) {
    if (mant == 0) return (uint16_t)(sign | 0x7C00);
    return (uint16_t)(sign | 0x7C00 | (mant >> 13));
  }
  int e = (int)exponent - 127 + 15;
  if (e <= 0) return (uint16_t)sign;
  if (e >= 0x1F) return (uint16_t)(sign | 0x7C00);
  return (uint16_t)(sign | ((uint32_t)e << 10) | (mant >> 13));
}
/* float16.go:21-50 ToFloat32 */
float orc_f16_to_f32(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000) << 16;
  uint32_t exponent = (h & 0x7C00) >> 10, mant = h & 0x3FF;
  if (exponent == 0) {
    if (mant == 0) return u2f(sign);
    uint32_t e = 1;
    while ((mant & 0x200) == 0) { mant <<= 1; e++; }
    mant &= 0x1FF;
    uint32_t ebits = (uint16_t)(127 - 15 - e + 1);
    return u2f(sign | (ebits << 23) | (mant << 13));
  }
  if (exponent == 0x1F) {
    if (mant == 0) return u2f(sign | 0x7F800000);
    return u2f(sign | 0x7FC00000 | (mant << 13));
  }
  return u2f(sign | ((exponent + 127 - 15) << 23) | (mant << 13));
}

/* bfloat16.go:104-167 GEMMBFloat16: ignores transA/transB; per element
 * sum = beta*c (or 0); for kk ascending: sum += (alpha*a)*b (all unfused fp32); store ToBFloat16(sum). */
void orc_gemm_bf16(int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb,
                   float beta, uint16_t* c, int ldc) {
  for (int i = 0; i < m; i++)
    for (int j = 0; j < n; j++) {
      float sum = 0.0f;
      if (beta != 0) sum = beta * orc_bf16_to_f32(c[(size_t)i * ldc + j]);
      for (int kk = 0; kk < k; kk++) {
        float t = alpha * orc_bf16_to_f32(a[(size_t)i * lda + kk]);
        t = t * orc_bf16_to_f32(b[(size_t)kk * ldb + j]);
        sum = sum + t;
      }
      c[(size_t)i * ldc + j] = orc_to_bf16(sum);
    }
}

/* float16.go:201-244 GEMMFloat16: upconvert the m*k / k*n / m*n element blocks,
 * fp32 guda.GEMM (with its fast paths), truncate back.  The reference converts
 * contiguous blocks, i.e. it is only meaningful for packed operands. */
int orc_gemm_f16(int ta, int tb, int m, int n, int k, float alpha, const uint16_t* a, int lda,
                 const uint16_t* b, int ldb, float beta, uint16_t* c, int ldc, int threads) {
  size_t as = (size_t)m * k, bs = (size_t)k * n, cs = (size_t)m * n;
  float* af = (float*)malloc((as + 1) * 4); float* bf = (float*)malloc((bs + 1) * 4); float* cf = (float*)calloc(cs + 1, 4);
  for (size_t i = 0; i < as; i++) af[i] = orc_f16_to_f32(a[i]);
  for (size_t i = 0; i < bs; i++) bf[i] = orc_f16_to_f32(b[i]);
  if (beta != 0) for (size_t i = 0; i < cs; i++) cf[i] = orc_f16_to_f32(c[i]);
  int rc = orc_sgemm(ta, tb, m, n, k, alpha, af, as, lda, bf, bs, ldb, beta, cf, cs, ldc, 0, threads);
  if (rc == ORC_OK) for (size_t i = 0; i < cs; i++) c[i] = orc_f32_to_f16(cf[i]);
  free(af); free(bf); free(cf);
  return rc;
}

/* ---------------------------------------------------------------- generators */

/* test_data.go:16-24 as written: 64-bit LCG state, float32(rng)/2^32 (NOT in [0,1)) */
void orc_generate_f32(float* out, size_t n, uint64_t seed) {
  uint64_t rng = seed;
  for (size_t i = 0; i < n; i++) {
    rng = rng * 1103515245ULL + 12345ULL;
    out[i] = (float)rng / 4294967296.0f;
  }
}
/* same LCG with the state masked to 32 bits; top 24 bits -> u in [0,1); u*(hi-lo)+lo as
 * GenerateFloat32Range does (test_data.go:37-44), unfused */
void orc_generate_f32_masked(float* out, size_t n, uint64_t seed, float lo, float hi) {
  uint32_t s = (uint32_t)seed;
  float scale = hi - lo;
  for (size_t i = 0; i < n; i++) {
    s = s * 1103515245u + 12345u;
    float u = (float)(s >> 8) * (1.0f / 16777216.0f);
    float v = u * scale;
    out[i] = v + lo;
  }
}
void orc_generate_f64_masked(double* out, size_t n, uint64_t seed, double lo, double hi) {
  uint32_t s = (uint32_t)seed;
  double scale = hi - lo;
  for (size_t i = 0; i < n; i++) {
    s = s * 1103515245u + 12345u;
    double u = (double)s * (1.0 / 4294967296.0);
    double v = u * scale;
    out[i] = v + lo;
  }
}

/* ---------------------------------------------------------------- tolerance */

/* tolerance.go:100-118 Float32ULPDiff */
int orc_f32_ulp_diff(float a, float b) {
  uint32_t ab = f2u(a), bb = f2u(b);
  if ((ab ^ bb) & 0x80000000u) return 0x7FFFFFFF;
  uint32_t d = ab > bb ? ab - bb : bb - ab;
  return (int)d;
}
/* tolerance.go:57-97 Float32NearEqual */
int orc_f32_near_equal(float a, float b, const orc_tol_t* tol) {
  if (tol->check_nan && isnan(a) && isnan(b)) return 1;
  if (tol->check_inf) {
    if (isinf(a) && a > 0 && isinf(b) && b > 0) return 1;
    if (isinf(a) && a < 0 && isinf(b) && b < 0) return 1;
  }
  if (a == b) return 1;
  float d32 = a - b;
  double diff = fabs((double)d32);
  if (diff <= (double)tol->abs_tol) return 1;
  double larger = fmax(fabs((double)a), fabs((double)b));
  if (diff <= larger * (double)tol->rel_tol) return 1;
  if (tol->ulp_tol > 0 && orc_f32_ulp_diff(a, b) <= tol->ulp_tol) return 1;
  return 0;
}
/* tolerance.go:131-174 VerifyFloat32Array */
void orc_verify_f32(const float* expected, const float* actual, size_t n, const orc_tol_t* tol, orc_verify_t* out) {
  out->max_abs = 0; out->max_rel = 0; out->max_ulp = 0; out->num_errors = 0; out->total = (int64_t)n; out->first_error = -1;
  for (size_t i = 0; i < n; i++) {
    if (orc_f32_near_equal(expected[i], actual[i], tol)) continue;
    out->num_errors++;
    if (out->first_error == -1) out->first_error = (int64_t)i;
    float ad = (float)fabs((double)(expected[i] - actual[i]));
    if (ad > out->max_abs) out->max_abs = ad;
    if (expected[i] != 0) { float rd = ad / (float)fabs((double)expected[i]); if (rd > out->max_rel) out->max_rel = rd; }
    int ud = orc_f32_ulp_diff(expected[i], actual[i]);
    if (ud > out->max_ulp) out->max_ulp = ud;
  }
}
