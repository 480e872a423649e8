/* level2.c -- CPU ORACLE (TEST INFRASTRUCTURE ONLY) for SGEMV / SGER: a restatement of
 *   compute/level2float32.go:244-308 (Sgemv argument checks and quick returns), :24-65 (Sger)
 *   compute/asm/f32/gemv.go:12-45 (GemvN), :52-91 (GemvT), compute/asm/f32/ge_amd64.s (Ger)
 * on the AVX2+FMA route: GemvN = one DotUnitary (32-lane blocked FMA order, dotunitary_avx2_amd64.s) or DotInc per row and
 * the UNFUSED combination y*beta + alpha*dot (plain Go arithmetic on amd64); GemvT = beta pass (zero fill / plain multiply)
 * then per row AxpyUnitaryTo (axpyunitaryto_amd64.s: MULPS + ADDPS, unfused) or AxpyInc (unfused); Ger = ge_amd64.s, whose 4x8 ... 1x1 blocks
 * all compute (alpha*x_i)*y_j with MULPS/MULSS and add a_ij with ADDPS/ADDSS (unfused).  Row-major, lda = row stride.
 * GemvT quirk kept: beta == 0 with incY == 1 clears the WHOLE y slice (gemv.go:61-64 `for i := range y`), not just n.
 * Negative increments follow the pure-Go twins (ge_noasm.go; kx = -(m-1)*incX): ge_amd64.s's prologue for incX/incY < 0 is
 * not exercised by any reference test (ge_test.go uses inc 1 and 2). */
#include "guda_oracle.h"

static int imax2(int a, int b) { return a > b ? a : b; }

/* compute/asm/f32/gemv.go:12-45 (GemvN) and :52-91 (GemvT) as exported by package f32: no argument checks and no alpha == 0
 * shortcut (0 * NaN stays NaN) -- this is the level gemv_test.go's golden vectors are stated at. */
void orc_f32_gemv(int trans, int m, int n, float alpha, const float* a, int lda, const float* x, int incx,
                  float beta, float* y, size_t leny_, int incy) {
  int lenx = trans ? m : n, leny = trans ? n : m;
  long long kx = incx < 0 ? -(long long)(lenx - 1) * incx : 0, ky = incy < 0 ? -(long long)(leny - 1) * incy : 0;
  if (!trans) {
    long long iy = ky;
    for (int i = 0; i < m; i++, iy += incy) {
      const float* row = a + (size_t)lda * i;
      float d = (incx == 1 && incy == 1) ? orc_sdot_unitary_avx2(row, x, (size_t)n) : orc_sdot_inc(x, row, (size_t)n, incx, 1, (size_t)kx, 0);
      float ad = alpha * d;
      if (beta == 0) y[iy] = ad;
      else { float yb = y[iy] * beta; y[iy] = yb + ad; }
    }
    return;
  }
  if (beta == 0) {
    if (incy == 1) for (size_t j = 0; j < leny_; j++) y[j] = 0.0f;                            /* whole slice: gemv.go:61-64 */
    else { long long iy = ky; for (int j = 0; j < n; j++, iy += incy) y[iy] = 0.0f; }
  }
  else { long long s = incy < 0 ? -incy : incy; for (int j = 0; j < n; j++) y[(long long)j * s] = y[(long long)j * s] * beta; }
  long long ix = kx;
  for (int i = 0; i < m; i++, ix += incx) {
    const float* row = a + (size_t)lda * i;
    float t = alpha * x[ix];
    if (incx == 1 && incy == 1) { for (int j = 0; j < n; j++) { float p = t * row[j]; y[j] = p + y[j]; } }   /* AxpyUnitaryTo */
    else orc_saxpy_inc(t, row, y, (size_t)n, 1, incy, 0, (size_t)ky);
  }
}

int orc_sgemv(int trans, int m, int n, float alpha, const float* a, size_t lena, int lda, const float* x, size_t lenx_, int incx,
              float beta, float* y, size_t leny_, int incy) {
  if (m < 0) return ORC_M_LT0;
  if (n < 0) return ORC_N_LT0;
  if (lda < imax2(1, n)) return ORC_BAD_LDA;
  if (incx == 0) return ORC_ZERO_INCX;
  if (incy == 0) return ORC_ZERO_INCY;
  int lenx = trans ? m : n, leny = trans ? n : m;
  if (m == 0 || n == 0) return ORC_OK;
  if ((incx > 0 && (long long)(lenx - 1) * incx >= (long long)lenx_) || (incx < 0 && (long long)(1 - lenx) * incx >= (long long)lenx_)) return ORC_SHORT_X;
  if ((incy > 0 && (long long)(leny - 1) * incy >= (long long)leny_) || (incy < 0 && (long long)(1 - leny) * incy >= (long long)leny_)) return ORC_SHORT_Y;
  if ((long long)lena < (long long)lda * (m - 1) + n) return ORC_SHORT_A;
  if (alpha == 0 && beta == 1) return ORC_OK;
  if (alpha == 0) return orc_sscal(leny, beta, y, leny_, incy > 0 ? incy : -incy);          /* level2float32.go:291-299 */
  orc_f32_gemv(trans, m, n, alpha, a, lda, x, incx, beta, y, leny_, incy);
  return ORC_OK;
}

int orc_sger(int m, int n, float alpha, const float* x, size_t lenx, int incx, const float* y, size_t leny, int incy,
             float* a, size_t lena, int lda) {
  if (m < 0) return ORC_M_LT0;
  if (n < 0) return ORC_N_LT0;
  if (lda < imax2(1, n)) return ORC_BAD_LDA;
  if (incx == 0) return ORC_ZERO_INCX;
  if (incy == 0) return ORC_ZERO_INCY;
  if (m == 0 || n == 0) return ORC_OK;
  if ((incx > 0 && (long long)lenx <= (long long)(m - 1) * incx) || (incx < 0 && (long long)lenx <= (long long)(1 - m) * incx)) return ORC_SHORT_X;
  if ((incy > 0 && (long long)leny <= (long long)(n - 1) * incy) || (incy < 0 && (long long)leny <= (long long)(1 - n) * incy)) return ORC_SHORT_Y;
  if ((long long)lena < (long long)lda * (m - 1) + n) return ORC_SHORT_A;
  if (alpha == 0) return ORC_OK;
  long long kx = incx < 0 ? -(long long)(m - 1) * incx : 0, ky = incy < 0 ? -(long long)(n - 1) * incy : 0;
  long long ix = kx;
  for (int i = 0; i < m; i++, ix += incx) {
    float t = alpha * x[ix];
    float* row = a + (size_t)lda * i;
    long long jy = ky;                                   /* ge_amd64.s: every block shape is MULPS/MULSS then ADDPS/ADDSS */
    for (int j = 0; j < n; j++, jy += incy) { float p = t * y[jy]; row[j] = p + row[j]; }
  }
  return ORC_OK;
}

/* ------------------------------------------------------------------ fp64: Dgemv / Dger
 * compute/level2float64.go:232-300 (Dgemv), :24-65 (Dger); kernels compute/asm/f64/gemvN_avx2_amd64.s (GemvNAVX2: per row 4 FMA
 * lanes over n/4 groups when incX == 1, (l0+l2)+(l1+l3), unfused remainder; incX != 1: one unfused scalar chain, then the same
 * fold with zero upper lanes; y_i = fl(alpha*s) + (beta == 0 ? 0 : fl(y_i*beta))), compute/asm/f64/gemvT_amd64.s (beta == 1: keep,
 * beta == 0: clear n, else scale; then per column the sequential unfused chain y_j += fl(fl(alpha*x_i) * a_ij) in row order) and
 * compute/asm/f64/ger_amd64.s (a_ij = fl(fl(alpha*x_i)*y_j) + a_ij).  Negative increments follow the BLAS definition. */
#include <math.h>
void orc_f64_gemv(int trans, int m, int n, double alpha, const double* a, int lda, const double* x, int incx,
                  double beta, double* y, int incy) {
  int lenx = trans ? m : n, leny = trans ? n : m;
  long long kx = incx < 0 ? -(long long)(lenx - 1) * incx : 0, ky = incy < 0 ? -(long long)(leny - 1) * incy : 0;
  if (!trans) {
    for (int i = 0; i < m; i++) { double* yp = y + ky + (long long)i * incy; *yp = beta == 0 ? 0.0 : *yp * beta; }
    for (int i = 0; i < m; i++) {
      const double* row = a + (size_t)lda * i;
      double l[4] = {0, 0, 0, 0};
      int j = 0;
      double s;
      if (incx == 1) {
        for (; j + 4 <= n; j += 4)
          for (int q = 0; q < 4; q++) l[q] = fma(row[j + q], x[j + q], l[q]);
        s = (l[0] + l[2]) + (l[1] + l[3]);
        for (; j < n; j++) { double p = row[j] * x[j]; s = s + p; }
      } else {
        const double* px = x + kx;
        for (; j < n; j++, px += incx) { double p = row[j] * px[0]; l[0] = l[0] + p; }
        s = (l[0] + 0.0) + (0.0 + 0.0);
      }
      double as = alpha * s;
      double* yp = y + ky + (long long)i * incy;
      *yp = as + *yp;
    }
    return;
  }
  if (beta == 0) { for (int j = 0; j < n; j++) y[ky + (long long)j * incy] = 0.0; }
  else if (beta != 1) { for (int j = 0; j < n; j++) { double* yp = y + ky + (long long)j * incy; *yp = *yp * beta; } }
  for (int i = 0; i < m; i++) {
    const double* row = a + (size_t)lda * i;
    double t = alpha * x[kx + (long long)i * incx];
    for (int j = 0; j < n; j++) { double p = t * row[j]; double* yp = y + ky + (long long)j * incy; *yp = *yp + p; }
  }
}

int orc_dgemv(int trans, int m, int n, double alpha, const double* a, size_t lena, int lda, const double* x, size_t lenx_, int incx,
              double beta, double* y, size_t leny_, int incy) {
  if (m < 0) return ORC_M_LT0;
  if (n < 0) return ORC_N_LT0;
  if (lda < imax2(1, n)) return ORC_BAD_LDA;
  if (incx == 0) return ORC_ZERO_INCX;
  if (incy == 0) return ORC_ZERO_INCY;
  int lenx = trans ? m : n, leny = trans ? n : m;
  if (m == 0 || n == 0) return ORC_OK;
  if ((incx > 0 && (long long)(lenx - 1) * incx >= (long long)lenx_) || (incx < 0 && (long long)(1 - lenx) * incx >= (long long)lenx_)) return ORC_SHORT_X;
  if ((incy > 0 && (long long)(leny - 1) * incy >= (long long)leny_) || (incy < 0 && (long long)(1 - leny) * incy >= (long long)leny_)) return ORC_SHORT_Y;
  if ((long long)lena < (long long)lda * (m - 1) + n) return ORC_SHORT_A;
  if (alpha == 0 && beta == 1) return ORC_OK;
  if (alpha == 0) return orc_dscal(leny, beta, y, leny_, incy > 0 ? incy : -incy);
  orc_f64_gemv(trans, m, n, alpha, a, lda, x, incx, beta, y, incy);
  return ORC_OK;
}

int orc_dger(int m, int n, double alpha, const double* x, size_t lenx, int incx, const double* y, size_t leny, int incy,
             double* a, size_t lena, int lda) {
  if (m < 0) return ORC_M_LT0;
  if (n < 0) return ORC_N_LT0;
  if (lda < imax2(1, n)) return ORC_BAD_LDA;
  if (incx == 0) return ORC_ZERO_INCX;
  if (incy == 0) return ORC_ZERO_INCY;
  if (m == 0 || n == 0) return ORC_OK;
  if ((incx > 0 && (long long)lenx <= (long long)(m - 1) * incx) || (incx < 0 && (long long)lenx <= (long long)(1 - m) * incx)) return ORC_SHORT_X;
  if ((incy > 0 && (long long)leny <= (long long)(n - 1) * incy) || (incy < 0 && (long long)leny <= (long long)(1 - n) * incy)) return ORC_SHORT_Y;
  if ((long long)lena < (long long)lda * (m - 1) + n) return ORC_SHORT_A;
  if (alpha == 0) return ORC_OK;
  long long kx = incx < 0 ? -(long long)(m - 1) * incx : 0, ky = incy < 0 ? -(long long)(n - 1) * incy : 0;
  for (int i = 0; i < m; i++) {
    double t = alpha * x[kx + (long long)i * incx];
    double* row = a + (size_t)lda * i;
    for (int j = 0; j < n; j++) { double p = t * y[ky + (long long)j * incy]; row[j] = p + row[j]; }
  }
  return ORC_OK;
}
