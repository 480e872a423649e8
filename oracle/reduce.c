/* reduce.c -- CPU ORACLE (TEST INFRASTRUCTURE ONLY) for the reductions of reduce.go:13-56 (DevicePtr.Sum/Max/Min/ArgMax/ArgMin/
 * SumSquares) and Isamax (compute/level1float32.go:89-130), restating the amd64 kernels they reach:
 *   compute/asm/f32/sum_amd64.s                  Sum: 16-byte alignment peel, 4x4 SSE lanes, fixed fold order
 *   compute/asm/f32/reduce_avx2_amd64.s:9-60     MaxAVX2  (VMAXPS: dst = src1 > src2 ? src1 : src2 -- NaN / +-0 return the NEW value)
 *   compute/asm/f32/reduce_avx2_amd64.s:62-113   MinAVX2
 *   compute/asm/f32/reduce_avx2_amd64.s:115-150  ArgMaxAVX2 (scalar loop, VCOMISS + JBE: update only on ordered strict >)
 *   compute/asm/f32/reduce_avx2_amd64.s:152-191  SumSquaresAVX2 (8 FMA lanes, lower+upper, two HADDs, FMA scalar tail)
 *   compute/asm/f32/reduce_avx2_amd64.s:193-228  ArgMinAVX2
 * Index and max/min outputs are what the GPU must match bit for bit; Sum / SumSquares are matched under ReduceArchTolerance. */
#include <math.h>
#include "guda_oracle.h"

static inline float op_max(float a, float b) { return a > b ? a : b; }   /* MAXPS/MAXSS: second operand unless a > b */
static inline float op_min(float a, float b) { return a < b ? a : b; }

static float minmax(const float* x, size_t n, int is_max) {
  if (n == 0) return is_max ? -INFINITY : INFINITY;
  float acc[8];
  for (int l = 0; l < 8; l++) acc[l] = x[0];
  size_t i = 0, rem = n;
  float r = x[0];
  if (n >= 8) {
    size_t groups = n >> 3;
    rem = n & 7;
    for (size_t g = 0; g < groups; g++, i += 8)
      for (int l = 0; l < 8; l++) acc[l] = is_max ? op_max(acc[l], x[i + l]) : op_min(acc[l], x[i + l]);
    float q[4];
    for (int l = 0; l < 4; l++) q[l] = is_max ? op_max(acc[l], acc[l + 4]) : op_min(acc[l], acc[l + 4]);
    for (int l = 0; l < 2; l++) q[l] = is_max ? op_max(q[l], q[l + 2]) : op_min(q[l], q[l + 2]);
    r = is_max ? op_max(q[0], q[1]) : op_min(q[0], q[1]);
  }
  for (; rem; rem--, i++) r = is_max ? op_max(r, x[i]) : op_min(r, x[i]);
  return r;
}
float orc_smax(const float* x, size_t n) { return minmax(x, n, 1); }
float orc_smin(const float* x, size_t n) { return minmax(x, n, 0); }

long long orc_sargmax(const float* x, size_t n) {
  if (n == 0) return -1;
  long long idx = 0;
  float best = x[0];
  for (size_t i = 1; i < n; i++)
    if (x[i] > best) { best = x[i]; idx = (long long)i; }       /* unordered (NaN on either side): no update */
  return idx;
}
long long orc_sargmin(const float* x, size_t n) {
  if (n == 0) return -1;
  long long idx = 0;
  float best = x[0];
  for (size_t i = 1; i < n; i++)
    if (best > x[i]) { best = x[i]; idx = (long long)i; }
  return idx;
}

/* compute/level1float32.go:89-130; *idx receives the result, the return value is the panic code */
int orc_isamax(int n, const float* x, size_t lenx, int incx, long long* idx) {
  *idx = -1;
  if (incx < 1) return incx == 0 ? ORC_ZERO_INCX : ORC_OK;
  if ((long long)lenx <= (long long)(n - 1) * incx) return ORC_SHORT_X;
  if (n < 2) {
    if (n == 1) { *idx = 0; return ORC_OK; }
    return n == 0 ? ORC_OK : ORC_N_LT0;
  }
  long long best = 0;
  float mx = fabsf(x[0]);
  for (int i = (incx == 1 ? 0 : 1); i < n; i++) {
    float a = fabsf(x[(long long)i * incx]);
    if (a > mx) { mx = a; best = i; }
  }
  *idx = best;
  return ORC_OK;
}

float orc_ssumsq(const float* x, size_t n) {
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  size_t i = 0, rem = n;
  float r = 0.0f;
  if (n >= 8) {
    size_t groups = n >> 3;
    rem = n & 7;
    for (size_t g = 0; g < groups; g++, i += 8)
      for (int l = 0; l < 8; l++) acc[l] = fmaf(x[i + l], x[i + l], acc[l]);
    float q[4];
    for (int l = 0; l < 4; l++) q[l] = acc[l] + acc[l + 4];      /* VADDPS X1(upper), X0(lower) */
    float h0 = q[0] + q[1], h1 = q[2] + q[3];                    /* VHADDPS */
    r = h0 + h1;                                                 /* VHADDPS */
  }
  for (; rem; rem--, i++) r = fmaf(x[i], x[i], r);
  return r;
}

/* sum_amd64.s: scalar adds into lane 0 until x is 16-byte aligned, then 16 lanes (4 SSE accumulators). */
float orc_ssum(const float* x, size_t n) {
  float s[4][4] = {{0}};
  if (n == 0) return 0.0f;
  size_t i = 0, len = n;
  while (((uintptr_t)(x + i) & 15) != 0) {
    s[0][0] = s[0][0] + x[i];
    i++; len--;
    if (len == 0) return s[0][0];
  }
  size_t tail = len;
  for (size_t it = 0; it < (len >> 4); it++, i += 16)
    for (int a = 0; a < 4; a++)
      for (int l = 0; l < 4; l++) s[a][l] = s[a][l] + x[i + 4 * a + l];
  for (int l = 0; l < 4; l++) { s[0][l] = s[0][l] + s[3][l]; s[1][l] = s[1][l] + s[2][l]; }
  if (tail & 8) {
    for (int l = 0; l < 4; l++) { s[0][l] = s[0][l] + x[i + l]; s[1][l] = s[1][l] + x[i + 4 + l]; }
    i += 8;
  }
  for (int l = 0; l < 4; l++) s[0][l] = s[0][l] + s[1][l];
  if (tail & 4) {
    for (int l = 0; l < 4; l++) s[0][l] = s[0][l] + x[i + l];
    i += 4;
  }
  float h0 = s[0][0] + s[0][1], h1 = s[0][2] + s[0][3];         /* HADDPS SUM, SUM -> [h0, h1, h0, h1] */
  if (tail & 2) { h0 = h0 + x[i]; h1 = h1 + x[i + 1]; i += 2; }
  float r = h0 + h1;                                             /* HADDPS */
  if (tail & 1) r = r + x[i];
  return r;
}

/* compute/asm/f32/reduce.go:152-186 fastExp: float32 arithmetic, unfused (Go does not contract on amd64) */
float orc_fast_exp(float x) {
  if (x < -87.3f) return 0.0f;
  const float ln2 = 0.6931472f, inv_ln2 = 1.4426951f;
  float s = x * inv_ln2;
  int32_t k = (int32_t)(x < 0 ? s - 0.5f : s + 0.5f);
  float kl = (float)k * ln2;
  float r = x - kl;
  const float c2 = 0.5f, c3 = 0.16666667f, c4 = 0.041666668f, c5 = 0.008333334f;
  float p = r * c5;  p = c4 + p;
  p = r * p;  p = c3 + p;
  p = r * p;  p = c2 + p;
  p = r * p;  p = 1.0f + p;
  p = r * p;  p = 1.0f + p;
  union { uint32_t u; float f; } sc;
  sc.u = (uint32_t)((k + 127) << 23);
  return p * sc.f;
}

/* reduce.go:180-187 -> compute/asm/f32/reduce.go:130-149 Softmax: Max (AVX2), fastExp(x - max), sequential sum, ScalUnitary(1/sum) */
void orc_softmax(float* x, size_t n) {
  if (n == 0) return;
  float mx = orc_smax(x, n);
  float sum = 0.0f;
  for (size_t i = 0; i < n; i++) { x[i] = orc_fast_exp(x[i] - mx); sum = sum + x[i]; }
  float inv = 1.0f / sum;
  for (size_t i = 0; i < n; i++) x[i] = x[i] * inv;
}

/* fused.go:270-310 LayerNorm: one pass sum / sum of squares (sequential, unfused), variance = E[x^2] - mean^2,
 * invStd = 1 / float32(sqrt(float64(variance + eps))), then (x - mean) * invStd [* gamma + beta when both are given] */
void orc_layer_norm(float* x, const float* gamma, const float* beta, size_t n, float eps) {
  float sum = 0.0f, sumsq = 0.0f;
  for (size_t i = 0; i < n; i++) { float v = x[i]; sum = sum + v; float q = v * v; sumsq = sumsq + q; }
  float mean = sum / (float)n;
  float ex2 = sumsq / (float)n, m2 = mean * mean;
  float variance = ex2 - m2;
  float ve = variance + eps;
  float inv = 1.0f / (float)sqrt((double)ve);
  for (size_t i = 0; i < n; i++) {
    float d = x[i] - mean;
    float nm = d * inv;
    if (gamma && beta) { float g = nm * gamma[i]; x[i] = g + beta[i]; }
    else x[i] = nm;
  }
}
