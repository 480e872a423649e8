/* gemm.c -- oracle (TEST INFRASTRUCTURE): SGEMM / DGEMM restatement.  See guda_oracle.h.
 *
 * SGEMM follows compute/sgemm.go on the AVX2+FMA route (tryOptimizedGemm returns
 * false, compute/sgemm_amd64.go:14-27 without AVX-512; the AVX-512 micro-kernel is
 * defective and is NOT a parity target -- SURVEY.md F8).  DGEMM follows
 * compute/dgemm.go -> dgemm_fix.go -> dgemmSerial (one thread). */
#include "guda_oracle.h"
#include <math.h>
#include <pthread.h>
#include <stdlib.h>

#define BLOCK 64        /* compute/gonum.go:20 */
#define MIN_PAR_BLOCK 4 /* compute/gonum.go:21 */

static inline int imax(int a, int b) { return a > b ? a : b; }

/* ---- 4^3 / 8^3 / 16^3 fast paths: compute/asm/f32/gemm_small.go:12-224 (unfused) ---- */
static inline float small_store(float alpha, float acc, float beta, float cold) {
  if (beta == 0) return alpha * acc;
  float p = alpha * acc, q = beta * cold;
  return p + q;
}
static void sgemm4x4(float alpha, const float* a, const float* b, float beta, float* c) {
  float r[16];
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      float p0 = a[i * 4 + 0] * b[0 * 4 + j], p1 = a[i * 4 + 1] * b[1 * 4 + j];
      float p2 = a[i * 4 + 2] * b[2 * 4 + j], p3 = a[i * 4 + 3] * b[3 * 4 + j];
      float s = p0 + p1; s = s + p2; s = s + p3;
      r[i * 4 + j] = s;
    }
  for (int i = 0; i < 16; i++) c[i] = small_store(alpha, r[i], beta, c[i]);
}
static void sgemm8x8(float alpha, const float* a, const float* b, float beta, float* c) {
  for (int i = 0; i < 8; i++)
    for (int j = 0; j < 8; j++) {
      float acc = 0.0f;
      for (int k = 0; k < 8; k += 2) {       /* c += a0*b0 + a1*b1 */
        float p0 = a[i * 8 + k] * b[k * 8 + j], p1 = a[i * 8 + k + 1] * b[(k + 1) * 8 + j];
        float s = p0 + p1;
        acc = acc + s;
      }
      c[i * 8 + j] = small_store(alpha, acc, beta, c[i * 8 + j]);
    }
}
static void sgemm16x16(float alpha, const float* a, const float* b, float beta, float* c) {
  for (int i = 0; i < 16; i++)
    for (int j = 0; j < 16; j++) {
      float acc = 0.0f;
      for (int k = 0; k < 16; k++) { float p = a[i * 16 + k] * b[k * 16 + j]; acc = acc + p; }
      c[i * 16 + j] = small_store(alpha, acc, beta, c[i * 16 + j]);
    }
}

/* ---- serial forms: compute/sgemm.go:288-350 ---- */
static void sgemm_serial(int ta, int tb, int m, int n, int k, const float* a, int lda,
                         const float* b, int ldb, float* c, int ldc, float alpha) {
  if (!ta && !tb) {                                   /* :297-305 */
    for (int i = 0; i < m; i++) {
      float* ctmp = c + (size_t)i * ldc;
      for (int l = 0; l < k; l++) {
        float tmp = alpha * a[(size_t)i * lda + l];
        if (tmp != 0) orc_saxpy_unitary_fma(tmp, b + (size_t)l * ldb, ctmp, (size_t)n);
      }
    }
  } else if (ta && !tb) {                             /* :309-321 */
    for (int l = 0; l < k; l++) {
      const float* btmp = b + (size_t)l * ldb;
      for (int i = 0; i < m; i++) {
        float tmp = alpha * a[(size_t)l * lda + i];
        if (tmp != 0) orc_saxpy_unitary_fma(tmp, btmp, c + (size_t)i * ldc, (size_t)n);
      }
    }
  } else if (!ta && tb) {                             /* :324-334 */
    for (int i = 0; i < m; i++) {
      const float* atmp = a + (size_t)i * lda;
      float* ctmp = c + (size_t)i * ldc;
      for (int j = 0; j < n; j++) {
        float d = alpha * orc_sdot_unitary_avx2(atmp, b + (size_t)j * ldb, (size_t)k);
        ctmp[j] = ctmp[j] + d;
      }
    }
  } else {                                            /* :337-349 */
    for (int l = 0; l < k; l++)
      for (int i = 0; i < m; i++) {
        float tmp = alpha * a[(size_t)l * lda + i];
        if (tmp != 0) orc_saxpy_inc(tmp, b + l, c + (size_t)i * ldc, (size_t)n, ldb, 1, 0, 0);
      }
  }
}

typedef struct {
  int ta, tb, m, n, k, lda, ldb, ldc, nbj, nblocks;
  const float *a, *b; float* c; float alpha;
  volatile int next;
} sg_job_t;

static void sgemm_block(const sg_job_t* J, int blk) {   /* compute/sgemm.go:216-262 */
  int i = (blk / J->nbj) * BLOCK, j = (blk % J->nbj) * BLOCK;
  int leni = i + BLOCK > J->m ? J->m - i : BLOCK;
  int lenj = j + BLOCK > J->n ? J->n - j : BLOCK;
  float* csub = J->c + (size_t)i * J->ldc + j;
  for (int kk = 0; kk < J->k; kk += BLOCK) {
    int lenk = kk + BLOCK > J->k ? J->k - kk : BLOCK;
    const float* asub = J->ta ? J->a + (size_t)kk * J->lda + i : J->a + (size_t)i * J->lda + kk;
    const float* bsub = J->tb ? J->b + (size_t)j * J->ldb + kk : J->b + (size_t)kk * J->ldb + j;
    sgemm_serial(J->ta, J->tb, leni, lenj, lenk, asub, J->lda, bsub, J->ldb, csub, J->ldc, J->alpha);
  }
}
static void* sgemm_worker(void* p) {
  sg_job_t* J = (sg_job_t*)p;
  for (;;) {
    int blk = __sync_fetch_and_add(&J->next, 1);
    if (blk >= J->nblocks) break;
    sgemm_block(J, blk);
  }
  return NULL;
}

int orc_sgemm(int ta, int tb, int m, int n, int k, float alpha,
              const float* a, size_t lena, int lda, const float* b, size_t lenb, int ldb,
              float beta, float* c, size_t lenc, int ldc, int stock, int threads) {
  /* compute/sgemm.go:30-71 */
  if (m < 0) return ORC_M_LT0;
  if (n < 0) return ORC_N_LT0;
  if (k < 0) return ORC_K_LT0;
  if (ta) { if (lda < imax(1, m)) return ORC_BAD_LDA; } else { if (lda < imax(1, k)) return ORC_BAD_LDA; }
  if (tb) { if (ldb < imax(1, k)) return ORC_BAD_LDB; } else { if (ldb < imax(1, n)) return ORC_BAD_LDB; }
  if (ldc < imax(1, n)) return ORC_BAD_LDC;
  if (m == 0 || n == 0) return ORC_OK;
  /* compute/sgemm.go:80-113: fast paths; note the bTrans branch calls the SAME
   * non-transposed kernels (a reference quirk, reproduced here because this is the oracle) */
  if (!stock && !ta && lda == k && ldc == n && m == n && n == k && (m == 4 || m == 8 || m == 16) &&
      ((!tb && ldb == n) || (tb && ldb == k))) {
    if (m == 4) sgemm4x4(alpha, a, b, beta, c);
    else if (m == 8) sgemm8x8(alpha, a, b, beta, c);
    else sgemm16x16(alpha, a, b, beta, c);
    return ORC_OK;
  }
  /* :116-136 */
  if (ta) { if ((long long)lena < (long long)(k - 1) * lda + m) return ORC_SHORT_A; }
  else    { if ((long long)lena < (long long)(m - 1) * lda + k) return ORC_SHORT_A; }
  if (tb) { if ((long long)lenb < (long long)(n - 1) * ldb + k) return ORC_SHORT_B; }
  else    { if ((long long)lenb < (long long)(k - 1) * ldb + n) return ORC_SHORT_B; }
  if ((long long)lenc < (long long)(m - 1) * ldc + n) return ORC_SHORT_C;
  if ((alpha == 0 || k == 0) && beta == 1) return ORC_OK;       /* :139 */
  if (beta != 1) {                                              /* :144-160 */
    if (beta == 0) { for (int i = 0; i < m; i++) for (int j = 0; j < n; j++) c[(size_t)i * ldc + j] = 0.0f; }
    else           { for (int i = 0; i < m; i++) for (int j = 0; j < n; j++) c[(size_t)i * ldc + j] = c[(size_t)i * ldc + j] * beta; }
  }
  /* sgemmParallel, compute/sgemm.go:165-264 */
  int nbi = (m + BLOCK - 1) / BLOCK, nbj = (n + BLOCK - 1) / BLOCK;
  if (nbi * nbj < MIN_PAR_BLOCK) { sgemm_serial(ta, tb, m, n, k, a, lda, b, ldb, c, ldc, alpha); return ORC_OK; }
  sg_job_t J = { ta, tb, m, n, k, lda, ldb, ldc, nbj, nbi * nbj, a, b, c, alpha, 0 };
  if (threads <= 1) { for (int blk = 0; blk < J.nblocks; blk++) sgemm_block(&J, blk); return ORC_OK; }
  if (threads > 256) threads = 256;
  pthread_t th[256];
  int started = 0;
  for (int t = 0; t < threads - 1; t++) { if (pthread_create(&th[started], NULL, sgemm_worker, &J) == 0) started++; }
  sgemm_worker(&J);
  for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
  return ORC_OK;
}

/* ------------------------------------------------------------------ DGEMM */

/* f64.GemmKernel8x8 / 4x4 (compute/asm/f64/gemm_kernel_4x4_amd64.s:36-83, 8x8 likewise):
 * load the C tile, for kk ascending c = fma(a[i,kk], b[kk,j], c), store. */
static void dgemm_tile_fma(const double* a, const double* b, double* c, int k, int lda, int ldb, int ldc, int mr, int nr) {
  for (int i = 0; i < mr; i++)
    for (int j = 0; j < nr; j++) {
      double acc = c[(size_t)i * ldc + j];
      for (int kk = 0; kk < k; kk++) acc = fma(a[(size_t)i * lda + kk], b[(size_t)kk * ldb + j], acc);
      c[(size_t)i * ldc + j] = acc;
    }
}
static void dgemm_edge(const double* a, const double* b, double* c, int k, int lda, int ldb, int ldc, int i, int j) {
  double sum = 0.0;                                    /* compute/dgemm_fixed.go:29-37 (unfused) */
  for (int l = 0; l < k; l++) { double p = a[(size_t)i * lda + l] * b[(size_t)l * ldb + j]; sum = sum + p; }
  c[(size_t)i * ldc + j] = c[(size_t)i * ldc + j] + sum;
}
/* compute/dgemm_fixed.go:6-68 */
static void dgemm_nn_kernel_fixed(int m, int n, int k, const double* a, int lda, const double* b, int ldb, double* c, int ldc) {
  int mi8 = m / 8 * 8, ni8 = n / 8 * 8, mi4 = m / 4 * 4, ni4 = n / 4 * 4;
  for (int i = 0; i < mi8; i += 8) {
    for (int j = 0; j < ni8; j += 8) dgemm_tile_fma(a + (size_t)i * lda, b + j, c + (size_t)i * ldc + j, k, lda, ldb, ldc, 8, 8);
    for (int j = ni8; j < ni4; j += 4) {
      dgemm_tile_fma(a + (size_t)i * lda, b + j, c + (size_t)i * ldc + j, k, lda, ldb, ldc, 4, 4);
      dgemm_tile_fma(a + (size_t)(i + 4) * lda, b + j, c + (size_t)(i + 4) * ldc + j, k, lda, ldb, ldc, 4, 4);
    }
    for (int j = ni4; j < n; j++) for (int ii = i; ii < i + 8 && ii < m; ii++) dgemm_edge(a, b, c, k, lda, ldb, ldc, ii, j);
  }
  for (int i = mi8; i < mi4; i += 4) {
    for (int j = 0; j < ni4; j += 4) dgemm_tile_fma(a + (size_t)i * lda, b + j, c + (size_t)i * ldc + j, k, lda, ldb, ldc, 4, 4);
    for (int j = ni4; j < n; j++) for (int ii = i; ii < i + 4 && ii < m; ii++) dgemm_edge(a, b, c, k, lda, ldb, ldc, ii, j);
  }
  for (int i = mi4; i < m; i++) for (int j = 0; j < n; j++) dgemm_edge(a, b, c, k, lda, ldb, ldc, i, j);
}

int orc_dgemm(int ta, int tb, int m, int n, int k, double alpha,
              const double* a, size_t lena, int lda, const double* b, size_t lenb, int ldb,
              double beta, double* c, size_t lenc, int ldc) {
  /* compute/dgemm.go:26-96 */
  if (m < 0) return ORC_M_LT0;
  if (n < 0) return ORC_N_LT0;
  if (k < 0) return ORC_K_LT0;
  if (ta) { if (lda < imax(1, m)) return ORC_BAD_LDA; } else { if (lda < imax(1, k)) return ORC_BAD_LDA; }
  if (tb) { if (ldb < imax(1, k)) return ORC_BAD_LDB; } else { if (ldb < imax(1, n)) return ORC_BAD_LDB; }
  if (ldc < imax(1, n)) return ORC_BAD_LDC;
  if (m == 0 || n == 0) return ORC_OK;
  if (ta) { if ((long long)lena < (long long)(k - 1) * lda + m) return ORC_SHORT_A; }
  else    { if ((long long)lena < (long long)(m - 1) * lda + k) return ORC_SHORT_A; }
  if (tb) { if ((long long)lenb < (long long)(n - 1) * ldb + k) return ORC_SHORT_B; }
  else    { if ((long long)lenb < (long long)(k - 1) * ldb + n) return ORC_SHORT_B; }
  if ((long long)lenc < (long long)(m - 1) * ldc + n) return ORC_SHORT_C;
  if ((alpha == 0 || k == 0) && beta == 1) return ORC_OK;       /* :99 */
  if (beta != 1) {                                              /* :104-119 */
    if (beta == 0) { for (int i = 0; i < m; i++) for (int j = 0; j < n; j++) c[(size_t)i * ldc + j] = 0.0; }
    else           { for (int i = 0; i < m; i++) for (int j = 0; j < n; j++) c[(size_t)i * ldc + j] = c[(size_t)i * ldc + j] * beta; }
  }
  /* dgemmParallelFixed -> dgemmSerial: compute/dgemm_fix.go:7-11, compute/dgemm.go:221-362 */
  if (!ta && !tb) {
    if (alpha == 1.0) { dgemm_nn_kernel_fixed(m, n, k, a, lda, b, ldb, c, ldc); return ORC_OK; }   /* :243-247 */
    for (int i = 0; i < m; i++)                                                                  /* :252-260 */
      for (int l = 0; l < k; l++) {
        double tmp = alpha * a[(size_t)i * lda + l];
        if (tmp != 0) orc_daxpy_unitary_fma(tmp, b + (size_t)l * ldb, c + (size_t)i * ldc, (size_t)n);
      }
  } else if (ta && !tb) {                                                                        /* :321-333 */
    for (int l = 0; l < k; l++)
      for (int i = 0; i < m; i++) {
        double tmp = alpha * a[(size_t)l * lda + i];
        if (tmp != 0) orc_daxpy_unitary_fma(tmp, b + (size_t)l * ldb, c + (size_t)i * ldc, (size_t)n);
      }
  } else if (!ta && tb) {                                                                        /* :336-346 */
    for (int i = 0; i < m; i++)
      for (int j = 0; j < n; j++) {
        double d = alpha * orc_ddot_unitary_avx2(a + (size_t)i * lda, b + (size_t)j * ldb, (size_t)k);
        c[(size_t)i * ldc + j] = c[(size_t)i * ldc + j] + d;
      }
  } else {                                                                                       /* :349-361 */
    for (int l = 0; l < k; l++)
      for (int i = 0; i < m; i++) {
        double tmp = alpha * a[(size_t)l * lda + i];
        if (tmp != 0) orc_daxpy_inc(tmp, b + l, c + (size_t)i * ldc, (size_t)n, ldb, 1, 0, 0);
      }
  }
  return ORC_OK;
}

/* ---- scalar Reference{}: reference.go:15-35, 86-145 (unfused; sum then c += alpha*sum) ---- */
void orc_ref_axpy(float alpha, const float* x, float* y, size_t n) {
  for (size_t i = 0; i < n; i++) { float p = alpha * x[i]; y[i] = p + y[i]; }
}
float orc_ref_dot(const float* x, const float* y, size_t n) {
  float sum = 0.0f;
  for (size_t i = 0; i < n; i++) { float p = x[i] * y[i]; sum = sum + p; }
  return sum;
}
void orc_ref_scale(float alpha, float* x, size_t n) { for (size_t i = 0; i < n; i++) x[i] = x[i] * alpha; }
void orc_ref_gemm(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda,
                  const float* b, int ldb, float beta, float* c, int ldc) {
  if (beta != 1.0f) for (int i = 0; i < m; i++) for (int j = 0; j < n; j++) c[(size_t)i * ldc + j] = c[(size_t)i * ldc + j] * beta;
  for (int i = 0; i < m; i++)
    for (int j = 0; j < n; j++) {
      float sum = 0.0f;
      for (int l = 0; l < k; l++) {
        float av = ta ? a[(size_t)l * lda + i] : a[(size_t)i * lda + l];
        float bv = tb ? b[(size_t)j * ldb + l] : b[(size_t)l * ldb + j];
        float p = av * bv;
        sum = sum + p;
      }
      float q = alpha * sum;
      c[(size_t)i * ldc + j] = c[(size_t)i * ldc + j] + q;
    }
}
