/* level1.c -- oracle (TEST INFRASTRUCTURE): Level-1 restatement.  See guda_oracle.h. */
#include "guda_oracle.h"
#include <math.h>

const char* orc_panic_string(int code) {
  /* compute/errors.go:9-35 */
  switch (code) {
    case ORC_OK: return "";
    case ORC_ZERO_INCX: return "blas: zero x index increment";
    case ORC_ZERO_INCY: return "blas: zero y index increment";
    case ORC_N_LT0: return "blas: n < 0";
    case ORC_M_LT0: return "blas: m < 0";
    case ORC_K_LT0: return "blas: k < 0";
    case ORC_BAD_LDA: return "blas: bad leading dimension of A";
    case ORC_BAD_LDB: return "blas: bad leading dimension of B";
    case ORC_BAD_LDC: return "blas: bad leading dimension of C";
    case ORC_SHORT_X: return "blas: insufficient length of x";
    case ORC_SHORT_Y: return "blas: insufficient length of y";
    case ORC_SHORT_A: return "blas: insufficient length of a";
    case ORC_SHORT_B: return "blas: insufficient length of b";
    case ORC_SHORT_C: return "blas: insufficient length of c";
    case ORC_NOT_SUPPORTED: return "operation not supported";
  }
  return "unknown";
}

/* ------------------------------------------------------------------ fp32 */

/* compute/asm/f32/axpyunitary_fma_amd64.s:20-136 -- every lane (alignment peel
 * :31-41, 32-wide body :50-70, 16/8/4 tails, scalar tail :126-133) is a fused
 * multiply-add, so the result is order- and alignment-independent. */
void orc_saxpy_unitary_fma(float alpha, const float* x, float* y, size_t n) {
  for (size_t i = 0; i < n; i++) y[i] = fmaf(alpha, x[i], y[i]);
}

/* compute/asm/f32/axpyinc_amd64.s:35-75 -- MULSS then ADDSS: unfused. */
void orc_saxpy_inc(float alpha, const float* x, float* y, size_t n, ptrdiff_t incx, ptrdiff_t incy, size_t ix, size_t iy) {
  const float* px = x + ix;
  float* py = y + iy;
  for (size_t i = 0; i < n; i++) {
    float t = px[0] * alpha;
    py[0] = t + py[0];
    px += incx; py += incy;
  }
}

/* compute/asm/f32/dotunitary_avx2_amd64.s:25-134.
 * 4 YMM accumulators x 8 lanes, FMA, 32 floats/iter (:49-68); reduce
 * (acc0+acc1)+(acc2+acc3) (:71-73); 16- and 8-float tails FMA into acc0 (:76-95);
 * lo128+hi128 (:99-100); 4-float tail MULPS+ADDPS unfused (:105-110);
 * horizontal (x0+x2),(x1+x3) then [0]+[1] (:114-118); scalar tail MULSS,ADDSS (:124-130). */
float orc_sdot_unitary_avx2(const float* x, const float* y, size_t n) {
  if (n == 0) return 0.0f;
  float acc[4][8];
  for (int a = 0; a < 4; a++) for (int l = 0; l < 8; l++) acc[a][l] = 0.0f;
  size_t idx = 0, len = n >> 5, tail = n & 31;
  if (len) {
    for (size_t it = 0; it < len; it++, idx += 32)
      for (int a = 0; a < 4; a++)
        for (int l = 0; l < 8; l++)
          acc[a][l] = fmaf(x[idx + 8 * a + l], y[idx + 8 * a + l], acc[a][l]);
    for (int l = 0; l < 8; l++) {
      float s01 = acc[0][l] + acc[1][l];
      float s23 = acc[2][l] + acc[3][l];
      acc[0][l] = s01 + s23;
    }
  }
  if (tail >= 16) {
    for (int l = 0; l < 8; l++) acc[0][l] = fmaf(x[idx + l], y[idx + l], acc[0][l]);
    for (int l = 0; l < 8; l++) acc[0][l] = fmaf(x[idx + 8 + l], y[idx + 8 + l], acc[0][l]);
    idx += 16; tail -= 16;
  }
  if (tail >= 8) {
    for (int l = 0; l < 8; l++) acc[0][l] = fmaf(x[idx + l], y[idx + l], acc[0][l]);
    idx += 8; tail -= 8;
  }
  float x0[4];
  for (int l = 0; l < 4; l++) x0[l] = acc[0][4 + l] + acc[0][l];   /* VADDPS X1(hi), X0(lo) */
  if (tail >= 4) {
    for (int l = 0; l < 4; l++) { float p = x[idx + l] * y[idx + l]; x0[l] = x0[l] + p; }
    idx += 4; tail -= 4;
  }
  float h0 = x0[0] + x0[2];
  float h1 = x0[1] + x0[3];
  float sum = h1 + h0;            /* ADDSS X1(=h0), X0(=h1) */
  while (tail) {
    float p = x[idx] * y[idx];
    sum = sum + p;
    idx++; tail--;
  }
  return sum;
}

/* compute/asm/f32/dotinc_amd64.s:40-88 -- two scalar accumulators (even/odd
 * elements of each group of 4), unfused; SUM += P_SUM; then scalar tail. */
float orc_sdot_inc(const float* x, const float* y, size_t n, ptrdiff_t incx, ptrdiff_t incy, size_t ix, size_t iy) {
  float sum = 0.0f;
  if (n == 0) return sum;
  const float* px = x + ix;
  const float* py = y + iy;
  size_t len = n >> 2, tail = n & 3;
  if (len) {
    float psum = 0.0f;
    for (size_t it = 0; it < len; it++) {
      float p0 = px[0] * py[0];
      float p1 = px[incx] * py[incy];
      float p2 = px[2 * incx] * py[2 * incy];
      float p3 = px[3 * incx] * py[3 * incy];
      sum = sum + p0; psum = psum + p1; sum = sum + p2; psum = psum + p3;
      px += 4 * incx; py += 4 * incy;
    }
    sum = sum + psum;
  }
  while (tail--) {
    float p = px[0] * py[0];
    sum = sum + p;
    px += incx; py += incy;
  }
  return sum;
}

/* compute/level1float32.go:225-259 */
int orc_saxpy(int n, float alpha, const float* x, size_t lenx, int incx, float* y, size_t leny, int incy) {
  if (incx == 0) return ORC_ZERO_INCX;
  if (incy == 0) return ORC_ZERO_INCY;
  if (n < 1) return n == 0 ? ORC_OK : ORC_N_LT0;
  long long L;
  L = (long long)(n - 1) * incx; if ((incx > 0 && (long long)lenx <= L) || (incx < 0 && (long long)lenx <= -L)) return ORC_SHORT_X;
  L = (long long)(n - 1) * incy; if ((incy > 0 && (long long)leny <= L) || (incy < 0 && (long long)leny <= -L)) return ORC_SHORT_Y;
  if (alpha == 0) return ORC_OK;
  if (incx == 1 && incy == 1) { orc_saxpy_unitary_fma(alpha, x, y, (size_t)n); return ORC_OK; }
  size_t ix = 0, iy = 0;
  if (incx < 0) ix = (size_t)((long long)(-n + 1) * incx);
  if (incy < 0) iy = (size_t)((long long)(-n + 1) * incy);
  orc_saxpy_inc(alpha, x, y, (size_t)n, incx, incy, ix, iy);
  return ORC_OK;
}

/* compute/level1float32_sdot.go:18-54 */
int orc_sdot(int n, const float* x, size_t lenx, int incx, const float* y, size_t leny, int incy, float* result) {
  *result = 0.0f;
  if (incx == 0) return ORC_ZERO_INCX;
  if (incy == 0) return ORC_ZERO_INCY;
  if (n <= 0) return n == 0 ? ORC_OK : ORC_N_LT0;
  if (incx == 1 && incy == 1) {
    if (lenx < (size_t)n) return ORC_SHORT_X;
    if (leny < (size_t)n) return ORC_SHORT_Y;
    *result = orc_sdot_unitary_avx2(x, y, (size_t)n);
    return ORC_OK;
  }
  long long ix = 0, iy = 0;
  if (incx < 0) ix = (long long)(-n + 1) * incx;
  if (incy < 0) iy = (long long)(-n + 1) * incy;
  if (ix >= (long long)lenx || ix + (long long)(n - 1) * incx >= (long long)lenx) return ORC_SHORT_X;
  if (iy >= (long long)leny || iy + (long long)(n - 1) * incy >= (long long)leny) return ORC_SHORT_Y;
  *result = orc_sdot_inc(x, y, (size_t)n, incx, incy, (size_t)ix, (size_t)iy);
  return ORC_OK;
}

/* compute/level1float32.go:619-653; compute/asm/f32/scal.go:12-16 (plain Go loop) */
int orc_sscal(int n, float alpha, float* x, size_t lenx, int incx) {
  if (incx < 1) return incx == 0 ? ORC_ZERO_INCX : ORC_OK;
  if (n < 1) return n == 0 ? ORC_OK : ORC_N_LT0;
  if ((long long)(n - 1) * incx >= (long long)lenx) return ORC_SHORT_X;
  if (alpha == 0) {
    for (long long i = 0; i < n; i++) x[i * incx] = 0.0f;   /* exact zero fill: NaN/Inf cleared */
    return ORC_OK;
  }
  for (long long i = 0; i < n; i++) x[i * incx] = x[i * incx] * alpha;
  return ORC_OK;
}

/* ------------------------------------------------------------------ fp64 */

/* compute/asm/f64/axpyunitary_fma_amd64.s -- alignment peel (until y is 32-byte
 * aligned) is MULSD+ADDSD (unfused); 16-wide body and 4-wide tail are
 * VFMADD213PD (fused); 2-wide SSE tail and the last single are unfused. */
void orc_daxpy_unitary_fma(double alpha, const double* x, double* y, size_t n) {
  size_t i = 0, len = n;
  if (len == 0) return;
  while (((uintptr_t)(y + i) & 31) != 0) {
    double t = x[i] * alpha;
    y[i] = t + y[i];
    i++; len--;
    if (len == 0) return;
  }
  size_t tail = len & 15, body = len >> 4;
  for (size_t it = 0; it < body; it++)
    for (int l = 0; l < 16; l++, i++) y[i] = fma(x[i], alpha, y[i]);
  if (tail == 0) return;
  for (size_t it = 0; it < (tail >> 2); it++)
    for (int l = 0; l < 4; l++, i++) y[i] = fma(x[i], alpha, y[i]);
  if ((tail & 3) >> 1) {
    for (int l = 0; l < 2; l++, i++) { double t = x[i] * alpha; y[i] = t + y[i]; }
  }
  if (tail & 1) { double t = x[i] * alpha; y[i] = t + y[i]; }
}

/* compute/asm/f64/axpyinc_amd64.s -- MULSD/ADDSD, unfused (same shape as the f32 one). */
void orc_daxpy_inc(double alpha, const double* x, double* y, size_t n, ptrdiff_t incx, ptrdiff_t incy, size_t ix, size_t iy) {
  const double* px = x + ix;
  double* py = y + iy;
  for (size_t i = 0; i < n; i++) {
    double t = px[0] * alpha;
    py[0] = t + py[0];
    px += incx; py += incy;
  }
}

/* compute/asm/f64/dotunitary_avx2_amd64.s -- 4 YMM accumulators x 4 lanes, FMA,
 * 16 doubles/iter; (Y0+Y1)+(Y2+Y3); 4-wide FMA tail into Y0; 2-wide MULPD+ADDPD
 * into the low half; 1-wide MULSD+ADDSD; lo128+hi128; [0]+[1]. */
double orc_ddot_unitary_avx2(const double* x, const double* y, size_t n) {
  double acc[4][4];
  for (int a = 0; a < 4; a++) for (int l = 0; l < 4; l++) acc[a][l] = 0.0;
  if (n == 0) return 0.0;
  size_t i = 0;
  for (size_t it = 0; it < (n >> 4); it++, i += 16)
    for (int a = 0; a < 4; a++)
      for (int l = 0; l < 4; l++)
        acc[a][l] = fma(x[i + 4 * a + l], y[i + 4 * a + l], acc[a][l]);
  for (int l = 0; l < 4; l++) {
    double s01 = acc[0][l] + acc[1][l];
    double s23 = acc[2][l] + acc[3][l];
    acc[0][l] = s01 + s23;
  }
  for (size_t it = 0; it < ((n & 15) >> 2); it++, i += 4)
    for (int l = 0; l < 4; l++) acc[0][l] = fma(x[i + l], y[i + l], acc[0][l]);
  if ((n & 3) >> 1) {
    for (int l = 0; l < 2; l++) { double p = x[i + l] * y[i + l]; acc[0][l] = acc[0][l] + p; }
    i += 2;
  }
  if (n & 1) { double p = x[i] * y[i]; acc[0][0] = acc[0][0] + p; }
  double l0 = acc[0][0] + acc[0][2];
  double l1 = acc[0][1] + acc[0][3];
  return l0 + l1;
}

/* compute/asm/f64/dot_amd64.s:100-142 DotInc -- one XMM accumulator: the high lane
 * sums products of elements 0,2,4,.. (MOVHPD), the low lane 1,3,5,.. (MOVLPD),
 * MULPD+ADDPD unfused; an odd last element is MULSD+ADDSD into the low lane;
 * result = high + low. */
double orc_ddot_inc(const double* x, const double* y, size_t n, ptrdiff_t incx, ptrdiff_t incy, size_t ix, size_t iy) {
  double lo = 0.0, hi = 0.0;
  const double* px = x + ix;
  const double* py = y + iy;
  long long cnt = (long long)n - 2;
  while (cnt >= 0) {
    double ph = px[0] * py[0];
    double pl = px[incx] * py[incy];
    hi = hi + ph; lo = lo + pl;
    px += 2 * incx; py += 2 * incy;
    cnt -= 2;
  }
  cnt += 2;
  if (cnt > 0) { double p = px[0] * py[0]; lo = lo + p; }
  return hi + lo;
}

/* compute/level1float64.go:211-246 */
int orc_daxpy(int n, double alpha, const double* x, size_t lenx, int incx, double* y, size_t leny, int incy) {
  if (incx == 0) return ORC_ZERO_INCX;
  if (incy == 0) return ORC_ZERO_INCY;
  if (n < 1) return n == 0 ? ORC_OK : ORC_N_LT0;
  long long L;
  L = (long long)(n - 1) * incx; if ((incx > 0 && (long long)lenx <= L) || (incx < 0 && (long long)lenx <= -L)) return ORC_SHORT_X;
  L = (long long)(n - 1) * incy; if ((incy > 0 && (long long)leny <= L) || (incy < 0 && (long long)leny <= -L)) return ORC_SHORT_Y;
  if (alpha == 0) return ORC_OK;
  if (incx == 1 && incy == 1) { orc_daxpy_unitary_fma(alpha, x, y, (size_t)n); return ORC_OK; }
  size_t ix = 0, iy = 0;
  if (incx < 0) ix = (size_t)((long long)(-n + 1) * incx);
  if (incy < 0) iy = (size_t)((long long)(-n + 1) * incy);
  orc_daxpy_inc(alpha, x, y, (size_t)n, incx, incy, ix, iy);
  return ORC_OK;
}

/* compute/level1float64_ddot.go:14-50 */
int orc_ddot(int n, const double* x, size_t lenx, int incx, const double* y, size_t leny, int incy, double* result) {
  *result = 0.0;
  if (incx == 0) return ORC_ZERO_INCX;
  if (incy == 0) return ORC_ZERO_INCY;
  if (n <= 0) return n == 0 ? ORC_OK : ORC_N_LT0;
  if (incx == 1 && incy == 1) {
    if (lenx < (size_t)n) return ORC_SHORT_X;
    if (leny < (size_t)n) return ORC_SHORT_Y;
    *result = orc_ddot_unitary_avx2(x, y, (size_t)n);
    return ORC_OK;
  }
  long long ix = 0, iy = 0;
  if (incx < 0) ix = (long long)(-n + 1) * incx;
  if (incy < 0) iy = (long long)(-n + 1) * incy;
  if (ix >= (long long)lenx || ix + (long long)(n - 1) * incx >= (long long)lenx) return ORC_SHORT_X;
  if (iy >= (long long)leny || iy + (long long)(n - 1) * incy >= (long long)leny) return ORC_SHORT_Y;
  *result = orc_ddot_inc(x, y, (size_t)n, incx, incy, (size_t)ix, (size_t)iy);
  return ORC_OK;
}

/* compute/level1float64.go:595-631; scalunitary_avx2 / scalinc are plain multiplies */
int orc_dscal(int n, double alpha, double* x, size_t lenx, int incx) {
  if (incx < 1) return incx == 0 ? ORC_ZERO_INCX : ORC_OK;
  if (n < 1) return n == 0 ? ORC_OK : ORC_N_LT0;
  if ((long long)(n - 1) * incx >= (long long)lenx) return ORC_SHORT_X;
  if (alpha == 0) {
    for (long long i = 0; i < n; i++) x[i * incx] = 0.0;
    return ORC_OK;
  }
  for (long long i = 0; i < n; i++) x[i * incx] = x[i * incx] * alpha;
  return ORC_OK;
}

/* ------------------------------------------------------------------ Dsdot / Sdsdot (fp32 inputs, fp64 accumulation)
 * compute/level1float32_dsdot.go:18-54, level1float32_sdsdot.go:18-54; kernels compute/asm/f32/ddotunitary_amd64.s (peel until y
 * is 16-byte aligned, then two 2-lane accumulators over groups of 8, pair tail, single tail, HADDPD) and ddotinc_amd64.s (two
 * scalar accumulators over groups of 4, then a sequential tail).  Products of two floats are exact in double. */
static double ddot_unitary(const float* x, const float* y, size_t n) {
  double s0 = 0, s1 = 0, p0 = 0, p1 = 0;
  if (n == 0) return 0.0;
  size_t i = 0, len = n;
  while (((uintptr_t)(y + i) & 15) != 0) {
    s0 = s0 + (double)x[i] * (double)y[i];
    i++; len--;
    if (len == 0) return s0 + s1;
  }
  size_t tail = len & 7;
  for (size_t it = 0; it < (len >> 3); it++, i += 8) {
    s0 = s0 + (double)x[i] * (double)y[i];         s1 = s1 + (double)x[i + 1] * (double)y[i + 1];
    p0 = p0 + (double)x[i + 2] * (double)y[i + 2]; p1 = p1 + (double)x[i + 3] * (double)y[i + 3];
    s0 = s0 + (double)x[i + 4] * (double)y[i + 4]; s1 = s1 + (double)x[i + 5] * (double)y[i + 5];
    p0 = p0 + (double)x[i + 6] * (double)y[i + 6]; p1 = p1 + (double)x[i + 7] * (double)y[i + 7];
  }
  if (len >> 3) { s0 = s0 + p0; s1 = s1 + p1; }
  for (size_t it = 0; it < (tail >> 1); it++, i += 2) { s0 = s0 + (double)x[i] * (double)y[i]; s1 = s1 + (double)x[i + 1] * (double)y[i + 1]; }
  if (tail & 1) s0 = s0 + (double)x[i] * (double)y[i];
  return s0 + s1;
}
static double ddot_inc(const float* x, const float* y, size_t n, ptrdiff_t incx, ptrdiff_t incy, size_t ix, size_t iy) {
  double sum = 0, psum = 0;
  const float* px = x + ix; const float* py = y + iy;
  size_t len = n >> 2, tail = n & 3;
  for (size_t it = 0; it < len; it++) {
    sum = sum + (double)px[0] * (double)py[0];               psum = psum + (double)px[incx] * (double)py[incy];
    sum = sum + (double)px[2 * incx] * (double)py[2 * incy]; psum = psum + (double)px[3 * incx] * (double)py[3 * incy];
    px += 4 * incx; py += 4 * incy;
  }
  if (len) sum = sum + psum;
  for (; tail; tail--, px += incx, py += incy) sum = sum + (double)px[0] * (double)py[0];
  return sum;
}
int orc_dsdot(int n, const float* x, size_t lenx, int incx, const float* y, size_t leny, int incy, double* result) {
  *result = 0.0;
  if (incx == 0) return ORC_ZERO_INCX;
  if (incy == 0) return ORC_ZERO_INCY;
  if (n <= 0) return n == 0 ? ORC_OK : ORC_N_LT0;
  if (incx == 1 && incy == 1) {
    if (lenx < (size_t)n) return ORC_SHORT_X;
    if (leny < (size_t)n) return ORC_SHORT_Y;
    *result = ddot_unitary(x, y, (size_t)n);
    return ORC_OK;
  }
  long long ix = 0, iy = 0;
  if (incx < 0) ix = (long long)(-n + 1) * incx;
  if (incy < 0) iy = (long long)(-n + 1) * incy;
  if (ix >= (long long)lenx || ix + (long long)(n - 1) * incx >= (long long)lenx) return ORC_SHORT_X;
  if (iy >= (long long)leny || iy + (long long)(n - 1) * incy >= (long long)leny) return ORC_SHORT_Y;
  *result = ddot_inc(x, y, (size_t)n, incx, incy, (size_t)ix, (size_t)iy);
  return ORC_OK;
}
int orc_sdsdot(int n, float alpha, const float* x, size_t lenx, int incx, const float* y, size_t leny, int incy, float* result) {
  double d;
  int rc = orc_dsdot(n, x, lenx, incx, y, leny, incy, &d);
  *result = 0.0f;
  if (rc || n <= 0) return rc;                      /* n == 0 returns 0, not alpha (level1float32_sdsdot.go:25-28) */
  *result = alpha + (float)d;
  return ORC_OK;
}
