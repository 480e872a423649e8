/*
 * guda_oracle.h -- CPU restatement of GUDA's BLAS hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is the parity oracle: a plain-C restatement of the arithmetic of the
 * reference (LynnColeArt/guda, Go + Plan-9 AVX2/FMA assembly).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.  The product (libguda_b200.so) never links or calls it.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * the reference checkout).  Build with -ffp-contract=off: a fused multiply-add
 * appears ONLY where the code says fmaf()/fma(), which is exactly where the
 * reference assembly uses VFMADD*.
 *
 * Parity pinning: see oracle/README.md -- the oracle reproduces the reference's
 * recorded baseline_20250807_221713.json checksums bit-for-bit and all of
 * blas/testblas's Dgemm/Daxpy/Ddot/Dscal answer tables (tests/test_oracle_golden.py).
 */
#ifndef GUDA_ORACLE_H
#define GUDA_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes mirror the panics of compute/errors.go:9-35 (0 = ok) */
enum {
  ORC_OK = 0,
  ORC_ZERO_INCX = 1, ORC_ZERO_INCY = 2, ORC_N_LT0 = 3, ORC_M_LT0 = 4, ORC_K_LT0 = 5,
  ORC_BAD_LDA = 6, ORC_BAD_LDB = 7, ORC_BAD_LDC = 8,
  ORC_SHORT_X = 9, ORC_SHORT_Y = 10, ORC_SHORT_A = 11, ORC_SHORT_B = 12, ORC_SHORT_C = 13,
  ORC_NOT_SUPPORTED = 14,
  ORC_BAD_CONV = 15        /* ConvParams.Validate (conv.go:34-51) */
};
const char* orc_panic_string(int code);

/* ---- Level 1, fp32 (compute/level1float32.go, compute/asm/f32) ---- */
void  orc_saxpy_unitary_fma(float alpha, const float* x, float* y, size_t n);
void  orc_saxpy_inc(float alpha, const float* x, float* y, size_t n, ptrdiff_t incx, ptrdiff_t incy, size_t ix, size_t iy);
float orc_sdot_unitary_avx2(const float* x, const float* y, size_t n);
float orc_sdot_inc(const float* x, const float* y, size_t n, ptrdiff_t incx, ptrdiff_t incy, size_t ix, size_t iy);
/* full Implementation{}.Saxpy/Sdot/Sscal with argument checks; lenx/leny = slice lengths */
int   orc_saxpy(int n, float alpha, const float* x, size_t lenx, int incx, float* y, size_t leny, int incy);
int   orc_sdot(int n, const float* x, size_t lenx, int incx, const float* y, size_t leny, int incy, float* result);
int   orc_sscal(int n, float alpha, float* x, size_t lenx, int incx);
int   orc_dsdot(int n, const float* x, size_t lenx, int incx, const float* y, size_t leny, int incy, double* result);   /* level1float32_dsdot.go */
int   orc_sdsdot(int n, float alpha, const float* x, size_t lenx, int incx, const float* y, size_t leny, int incy, float* result);

/* ---- Level 1, fp64 (compute/level1float64.go, compute/asm/f64) ---- */
void   orc_daxpy_unitary_fma(double alpha, const double* x, double* y, size_t n);
void   orc_daxpy_inc(double alpha, const double* x, double* y, size_t n, ptrdiff_t incx, ptrdiff_t incy, size_t ix, size_t iy);
double orc_ddot_unitary_avx2(const double* x, const double* y, size_t n);
double orc_ddot_inc(const double* x, const double* y, size_t n, ptrdiff_t incx, ptrdiff_t incy, size_t ix, size_t iy);
int    orc_daxpy(int n, double alpha, const double* x, size_t lenx, int incx, double* y, size_t leny, int incy);
int    orc_ddot(int n, const double* x, size_t lenx, int incx, const double* y, size_t leny, int incy, double* result);
int    orc_dscal(int n, double alpha, double* x, size_t lenx, int incx);

/* ---- Level 3 (compute/sgemm.go, compute/dgemm.go, blas/gonum/sgemm.go) ----
 * threads: worker count used for the 64x64 block fan-out (GOMAXPROCS); results
 * do not depend on it.  stock != 0 follows blas/gonum/sgemm.go (no 4/8/16 fast paths). */
int orc_sgemm(int ta, int tb, int m, int n, int k, float alpha,
              const float* a, size_t lena, int lda, const float* b, size_t lenb, int ldb,
              float beta, float* c, size_t lenc, int ldc, int stock, int threads);
int orc_dgemm(int ta, int tb, int m, int n, int k, double alpha,
              const double* a, size_t lena, int lda, const double* b, size_t lenb, int ldb,
              double beta, double* c, size_t lenc, int ldc);

/* ---- scalar Reference{} (reference.go:15-35,86-145) ---- */
void  orc_ref_axpy(float alpha, const float* x, float* y, size_t n);
float orc_ref_dot(const float* x, const float* y, size_t n);
void  orc_ref_scale(float alpha, float* x, size_t n);
void  orc_ref_gemm(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda,
                   const float* b, int ldb, float beta, float* c, int ldc);
void  orc_ref_gelu(float* x, size_t n);                      /* reference.go:203-211 */

/* ---- activations (activations.go:31-90, fused_gelu.go:44-56, gelu_avx2_amd64.s:46-77) ---- */
float orc_exp_f32(float x);
float orc_tanh_f32(float x);
float orc_gelu_f32(float x);         /* scalar route: geluFloat32 */
float orc_gelu_pade(float x);        /* AVX2 route: GeluAVX2 clamped Pade tanh */

/* ---- fused GEMM + bias + activation (fused.go:313-368, fused_gelu.go:65-137) ----
 * gelu_route: 0 = geluFloat32 (pinned), 1 = GeluAVX2 Pade (secondary) */
int orc_fused_gemm_bias_relu(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda,
                             const float* b, int ldb, float beta, float* c, int ldc, const float* bias, int threads);
int orc_fused_gemm_bias_gelu(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda,
                             const float* b, int ldb, float beta, float* c, int ldc, const float* bias,
                             int gelu_route, int threads);
void orc_add_bias_gelu(float* x, const float* bias, int n, int bias_len);     /* fused_gelu.go:12-41 */

/* ---- Level 2 (compute/level2float32.go:244-308 Sgemv, :24-65 Sger; compute/asm/f32/gemv.go, ge_noasm.go); row-major ---- */
void orc_f32_gemv(int trans, int m, int n, float alpha, const float* a, int lda, const float* x, int incx,
                  float beta, float* y, size_t leny, int incy);   /* package f32's GemvN/GemvT: no checks, no alpha == 0 shortcut */
int orc_sgemv(int trans, int m, int n, float alpha, const float* a, size_t lena, int lda, const float* x, size_t lenx, int incx,
              float beta, float* y, size_t leny, int incy);
void orc_f64_gemv(int trans, int m, int n, double alpha, const double* a, int lda, const double* x, int incx, double beta, double* y, int incy);
int orc_dgemv(int trans, int m, int n, double alpha, const double* a, size_t lena, int lda, const double* x, size_t lenx, int incx,
              double beta, double* y, size_t leny, int incy);                       /* compute/level2float64.go:232-300 */
int orc_dger(int m, int n, double alpha, const double* x, size_t lenx, int incx, const double* y, size_t leny, int incy,
             double* a, size_t lena, int lda);                                      /* compute/level2float64.go:24-65 */
int orc_sger(int m, int n, float alpha, const float* x, size_t lenx, int incx, const float* y, size_t leny, int incy,
             float* a, size_t lena, int lda);

/* ---- Reductions (reduce.go:13-56 -> compute/asm/f32/{sum_amd64.s,reduce_avx2_amd64.s}; Isamax compute/level1float32.go:89-130) ---- */
float orc_ssum(const float* x, size_t n);
float orc_ssumsq(const float* x, size_t n);
float orc_smax(const float* x, size_t n);
float orc_smin(const float* x, size_t n);
long long orc_sargmax(const float* x, size_t n);
long long orc_sargmin(const float* x, size_t n);
int orc_isamax(int n, const float* x, size_t lenx, int incx, long long* idx);
float orc_fast_exp(float x);                         /* compute/asm/f32/reduce.go:152-186 */
void orc_softmax(float* x, size_t n);                /* reduce.go:180 -> compute/asm/f32/reduce.go:130-149 */
void orc_layer_norm(float* x, const float* gamma, const float* beta, size_t n, float eps);   /* fused.go:270-310 */

/* ---- Conv2D via im2col -> Sgemm (conv.go:69-226) and the direct form (conv.go:231-300); NCHW, kernel [OC][IC][KH][KW] ---- */
void orc_im2col(const float* input, int channels, int height, int width, int kh, int kw, int pad_h, int pad_w,
                int stride_h, int stride_w, int dil_h, int dil_w, float* output, size_t ld_out);
int orc_conv2d(const float* input, const float* kernel, const float* bias, float* output, int batch, int in_c, int in_h, int in_w,
               int out_c, int kh, int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w, int use_bias,
               int threads);
int orc_conv2d_direct(const float* input, const float* kernel, const float* bias, float* output, int batch, int in_c, int in_h,
                      int in_w, int out_c, int kh, int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                      int use_bias);

/* ---- bf16 / fp16 (bfloat16.go:12-29,104-167; float16.go:21-87,201-244) ---- */
uint16_t orc_to_bf16(float f);
float    orc_bf16_to_f32(uint16_t b);
uint16_t orc_f32_to_f16(float f);
float    orc_f16_to_f32(uint16_t h);
void orc_gemm_bf16(int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb,
                   float beta, uint16_t* c, int ldc);
int  orc_gemm_f16(int ta, int tb, int m, int n, int k, float alpha, const uint16_t* a, int lda,
                  const uint16_t* b, int ldb, float beta, uint16_t* c, int ldc, int threads);

/* ---- generators (test_data.go:16-24 and a masked 32-bit variant) ---- */
void orc_generate_f32(float* out, size_t n, uint64_t seed);                 /* as written in the reference */
void orc_generate_f32_masked(float* out, size_t n, uint64_t seed, float lo, float hi);   /* 32-bit LCG state -> [lo,hi) */
void orc_generate_f64_masked(double* out, size_t n, uint64_t seed, double lo, double hi);

/* ---- tolerance framework (tolerance.go:10-191, tolerance_arch.go:8-127) ---- */
/* ---- elementwise.c: math.go:133-169,230-272, activations.go:9-27, fused.go:18-216 ---- */
typedef struct { int type; float alpha; long long bsize; long long bdim; } orc_fused_op_t;
float orc_sigmoid_f32(float x);
void orc_ew_add(const float* a, const float* b, float* c, size_t n);
void orc_ew_mul(const float* a, const float* b, float* c, size_t n);
void orc_ew_relu(float* x, size_t n);
void orc_ew_sigmoid(float* x, size_t n);
int orc_reduce_axis(const float* x, int m, int n, int axis, int is_max, float* out);      /* reduce.go:105-178 */
int orc_topk(const float* x, long long n, int k, float* values, long long* indices);            /* reduce.go:247-291 */
void orc_fused_execute(float* x, const float* const* others, int n_others, size_t n, const orc_fused_op_t* ops, int n_ops);

typedef struct { float abs_tol, rel_tol; int ulp_tol; int check_nan, check_inf; } orc_tol_t;
int  orc_f32_ulp_diff(float a, float b);
int  orc_f32_near_equal(float a, float b, const orc_tol_t* tol);
typedef struct { float max_abs, max_rel; int max_ulp; int64_t num_errors, total, first_error; } orc_verify_t;
void orc_verify_f32(const float* expected, const float* actual, size_t n, const orc_tol_t* tol, orc_verify_t* out);

#ifdef __cplusplus
}
#endif
#endif
