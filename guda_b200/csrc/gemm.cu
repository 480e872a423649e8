// gemm.cu -- SGEMM / DGEMM / fused GEMM+bias+activation entry points on the exact-order SIMT kernels.
// Replaces compute/sgemm.go:29-350 (Sgemm), compute/dgemm.go:25-362 + dgemm_fixed.go (Dgemm),
// fused.go:313-368 and fused_gelu.go:65-137.
#include "gemm_simt.cuh"
#include <cstdlib>

namespace guda {

int sgemm_mode();
// FP64 tensor pipe (dgemm_dmma.cu); GUDA_OK / error if it took the call, -1 if the operands are not eligible
int dgemm_dmma_try(int ta, int tb, int m, int n, int k, double alpha, const double* a, int lda, const double* b, int ldb, double beta,
                   double* c, int ldc, int edge_m, int edge_n, bool skip, cudaStream_t s, const ScatterArgsT<double>* sc = nullptr);
// TMA-fed exact-order pipeline (sgemm_pipe.cu); GUDA_OK / error if it took the call, -1 if the call is not eligible
int sgemm_pipe_try(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb, float beta, float* c,
                   int ldc, const float* bias, int epi, const ScatterArgs* scatter, cudaStream_t s);
// tensor-core paths (gemm_tc.cu); return GUDA_OK if they handled the call, -1 if the shape is not eligible
int sgemm_3xtf32_try(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb,
                     float beta, float* c, int ldc, cudaStream_t s);

// ---------------------------------------------------------------- tiny fast paths (compute/asm/f32/gemm_small.go:12-224)
// 4^3 / 8^3 / 16^3 contiguous NN: unfused products summed in the reference's order, then alpha*s (+ beta*c).
template <int S>
__global__ void sgemm_small_kernel(float alpha, const float* __restrict__ a, const float* __restrict__ b, float beta, float* __restrict__ c) {
  int i = threadIdx.x / S, j = threadIdx.x % S;
  float s;
  if (S == 4) {
    s = __fadd_rn(__fmul_rn(a[i * 4 + 0], b[0 * 4 + j]), __fmul_rn(a[i * 4 + 1], b[1 * 4 + j]));
    s = __fadd_rn(s, __fmul_rn(a[i * 4 + 2], b[2 * 4 + j]));
    s = __fadd_rn(s, __fmul_rn(a[i * 4 + 3], b[3 * 4 + j]));
  } else if (S == 8) {
    s = 0.0f;
    for (int k = 0; k < 8; k += 2) {
      float p = __fadd_rn(__fmul_rn(a[i * 8 + k], b[k * 8 + j]), __fmul_rn(a[i * 8 + k + 1], b[(k + 1) * 8 + j]));
      s = __fadd_rn(s, p);
    }
  } else {
    s = 0.0f;
    for (int k = 0; k < 16; k++) s = __fadd_rn(s, __fmul_rn(a[i * 16 + k], b[k * 16 + j]));
  }
  float cold = c[i * S + j];
  __syncthreads();
  c[i * S + j] = (beta == 0.0f) ? __fmul_rn(alpha, s) : __fadd_rn(__fmul_rn(alpha, s), __fmul_rn(beta, cold));
}

// ---------------------------------------------------------------- fp64 NN alpha==1 edges (compute/dgemm_fixed.go:29-37,48-56,60-67)
// rows >= mi4 (all columns) and columns >= ni4 (rows < mi4): sum = 0; sum += a*b unfused; c = beta-scaled c + sum
__global__ void dgemm_edge_kernel(int m, int n, int k, const double* __restrict__ a, long long lda, const double* __restrict__ b,
                                  long long ldb, double beta, double* __restrict__ c, long long ldc, int mi4, int ni4) {
  long long nrow_edge = (long long)(m - mi4) * n;
  long long ncol_edge = (long long)mi4 * (n - ni4);
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < nrow_edge + ncol_edge; e += (long long)gridDim.x * blockDim.x) {
    int i, j;
    if (e < nrow_edge) { i = mi4 + (int)(e / n); j = (int)(e % n); }
    else { long long f = e - nrow_edge; j = ni4 + (int)(f / mi4); i = (int)(f % mi4); }    // consecutive threads walk down a column
    double sum = 0.0;
    for (int l = 0; l < k; l++) sum = __dadd_rn(sum, __dmul_rn(a[(long long)i * lda + l], b[(long long)l * ldb + j]));
    double cv = 0.0;
    if (beta != 0.0) { cv = c[(long long)i * ldc + j]; if (beta != 1.0) cv = __dmul_rn(cv, beta); }
    c[(long long)i * ldc + j] = __dadd_rn(cv, sum);
  }
}

// ---------------------------------------------------------------- dispatch
template <typename T> struct TileCfg;
template <> struct TileCfg<float>  { static constexpr int BK = 16, MINB_BIG = 2, MINB_SMALL = 3; };
template <> struct TileCfg<double> { static constexpr int BK = 8,  MINB_BIG = 1, MINB_SMALL = 2; };

template <typename T, bool TA, bool TB, int MODE, int EPI, int BM, int BN, int MINB, bool SCAT = false, int TXN = 16>
static int launch_simt_tile(const GemmParams<T>& p, cudaStream_t s, const ScatterParams<T, SCAT>& sp) {
  constexpr int BK = TileCfg<T>::BK;
  constexpr int VN = GemmVec<T>::N;
  auto aligned = [&](const void* ptr, long long ld) { return ((uintptr_t)ptr % 16 == 0) && (ld % VN == 0); };
  const bool fast = (p.m % BM == 0) && (p.n % BN == 0) && (p.k % BK == 0) && aligned(p.a, p.lda) && aligned(p.b, p.ldb) &&
                    aligned(p.c, p.ldc) && p.edge_m >= p.m && p.edge_n >= p.n;
  dim3 grid((p.n + BN - 1) / BN, (p.m + BM - 1) / BM);
  if (grid.y > 65535) return set_error(GUDA_ERR_INVALID_ARG, "m too large for one launch");
  if (fast) gemm_simt_kernel<T, TA, TB, BM, BN, BK, MODE, EPI, true, MINB, SCAT, TXN><<<grid, 256, 0, s>>>(p, sp);
  else      gemm_simt_kernel<T, TA, TB, BM, BN, BK, MODE, EPI, false, MINB, SCAT, TXN><<<grid, 256, 0, s>>>(p, sp);
  return check_launch("gemm");
}

// Tile choice: every C element is one thread's sequential-k chain, so there is no split-K to fill the machine with;
// the only freedom is the CTA tile.  Pick the shape with the smallest estimated makespan
//   ceil(tiles / SMs) * BM*BN / efficiency      (efficiency = FFMA share of the inner loop, measured: 8x8 micro-tile 1.0,
//   8x4 0.85-0.93, 4x4 0.6), which sends 1024^3 (64 big tiles on 148 SMs) to the middle tile and anything large to 128x128.
template <typename T, bool TA, bool TB, int MODE, int EPI, bool SCAT = false>
static int launch_simt_one(const GemmParams<T>& p, cudaStream_t s, const ScatterParams<T, SCAT>& sp = {}) {
  const long long sms = sm_count();
  auto makespan = [&](int bm, int bn, double eff) {
    const long long tiles = (long long)((p.m + bm - 1) / bm) * ((p.n + bn - 1) / bn);
    return (double)((tiles + sms - 1) / sms) * bm * bn / eff;
  };
  // fp32 FMA forms: the middle tile is 64 x 128 on a 32 x 8 thread grid (8x4 micro-tile; a warp is one m-row of threads, so its
  // A fragments are warp-wide broadcasts and its B fragments 8-byte loads): 6-8 shared-memory wavefronts per k step where the
  // 128 x 64 tile on the 16 x 16 grid needs 8 for the same 8 FMA cycles.  Measured 7 % faster (1024^3: 61.5 -> 57.4 us).
  constexpr bool WIDE_MID = sizeof(T) == 4 && !SCAT && MODE == MODE_FMA;
  double t_big = makespan(128, 128, 1.0), t_mid = WIDE_MID ? makespan(64, 128, 0.93) : makespan(128, 64, 0.85), t_small = makespan(64, 64, 0.6);
  static const int forced = [] { const char* e = getenv("GUDA_SIMT_TILE"); return e ? atoi(e) : 0; }();   // A/B knob: 1 big, 2 mid, 3 small, 4 legacy mid
  if (forced == 1) t_big = 0; else if (forced == 2 || forced == 4) { t_mid = 0; t_big = 1; t_small = 1; } else if (forced == 3) { t_small = 0; t_big = 1; t_mid = 1; }
  if (t_big <= t_mid && t_big <= t_small) return launch_simt_tile<T, TA, TB, MODE, EPI, 128, 128, TileCfg<T>::MINB_BIG, SCAT>(p, s, sp);
  if constexpr (sizeof(T) == 4) {
    if (t_mid <= t_small) {
      if constexpr (WIDE_MID) {
        if (forced != 4) return launch_simt_tile<T, TA, TB, MODE, EPI, 64, 128, 2, SCAT, 32>(p, s, sp);
      }
      return launch_simt_tile<T, TA, TB, MODE, EPI, 128, 64, 2, SCAT>(p, s, sp);
    }
  }
  return launch_simt_tile<T, TA, TB, MODE, EPI, 64, 64, TileCfg<T>::MINB_SMALL, SCAT>(p, s, sp);
}

// K blocking across launches.  A C element is one sequential-k chain held in an fp32 (fp64) accumulator, so stopping the
// chain at k1, storing the accumulator and resuming from it with beta = 1 gives the SAME bits (the reference itself walks
// 64-deep k blocks this way, compute/sgemm.go:236-260).  With K far beyond what L2 can hold, the ~300 co-resident CTAs
// drift apart in k and stop sharing A/B panels: ncu measured 11.4 TB of DRAM reads for one 65536^3 launch (L2 hit 35 %).
// Chunks of KC keep the resident set -- (16 + ~18) tile panels x KC -- inside L2.
// KC = 8192 (fp32): a launch boundary is a grid-wide barrier, so the drift is bounded by one chunk while the extra C
// round trip (2 x |C| per chunk) and the pipeline refill stay near 1 % of the work (KC = 2048 cut traffic to 1.7 TB but cost 5 %).
template <typename T> struct KChunk { static constexpr int KC = sizeof(T) == 4 ? 8192 : 4096; };

template <typename T, bool TA, bool TB, int MODE, int EPI, bool SCAT = false>
static int launch_simt(const GemmParams<T>& p, cudaStream_t s, const ScatterParams<T, SCAT>& sp = {}) {
  constexpr int KC = KChunk<T>::KC;
  const bool chunk = EPI == EPI_NONE && p.c != nullptr && p.k >= 2 * KC && (long long)p.m * p.n >= (1ll << 22) &&
                     p.edge_m >= p.m && p.edge_n >= p.n;
  if (!chunk) return launch_simt_one<T, TA, TB, MODE, EPI, SCAT>(p, s, sp);
  GemmParams<T> q = p;
  for (int k0 = 0; k0 < p.k; k0 += KC) {
    const bool last = k0 + KC >= p.k;
    q.k = last ? p.k - k0 : KC;
    q.a = TA ? p.a + (long long)k0 * p.lda : p.a + k0;
    q.b = TB ? p.b + k0 : p.b + (long long)k0 * p.ldb;
    q.beta = k0 == 0 ? p.beta : T(1);
    int rc = (SCAT && last) ? launch_simt_one<T, TA, TB, MODE, EPI, SCAT>(q, s, sp) : launch_simt_one<T, TA, TB, MODE, EPI, false>(q, s);
    if (rc) return rc;
  }
  return GUDA_OK;
}

// argument checks of compute/sgemm.go:30-71 (same order, same panic strings).  Slice-length checks
// (:116-136) need the DevicePtr sizes and live in the host shim.
static int check_gemm_args(int ta, int tb, int m, int n, int k, int lda, int ldb, int ldc) {
  if ((ta != 0 && ta != 1) || (tb != 0 && tb != 1)) GUDA_PANIC("blas: illegal transpose");
  if (m < 0) GUDA_PANIC("blas: m < 0");
  if (n < 0) GUDA_PANIC("blas: n < 0");
  if (k < 0) GUDA_PANIC("blas: k < 0");
  auto mx = [](int a, int b) { return a > b ? a : b; };
  if (ta) { if (lda < mx(1, m)) GUDA_PANIC("blas: bad leading dimension of A"); }
  else    { if (lda < mx(1, k)) GUDA_PANIC("blas: bad leading dimension of A"); }
  if (tb) { if (ldb < mx(1, k)) GUDA_PANIC("blas: bad leading dimension of B"); }
  else    { if (ldb < mx(1, n)) GUDA_PANIC("blas: bad leading dimension of B"); }
  if (ldc < mx(1, n)) GUDA_PANIC("blas: bad leading dimension of C");
  return GUDA_OK;
}

static int sgemm_impl(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb,
                      float beta, float* c, int ldc, const float* bias, int epi, cudaStream_t s,
                      const ScatterArgs* scatter = nullptr, bool force_exact = false) {
  int rc = check_gemm_args(ta, tb, m, n, k, lda, ldb, ldc);
  if (rc) return rc;
  if (m == 0 || n == 0) return GUDA_OK;                                   // compute/sgemm.go:74
  rc = bind_device();
  if (rc) return rc;
  if ((!c && !(scatter && beta == 0.0f)) || ((!a || !b) && k > 0)) return set_error(GUDA_ERR_INVALID_ARG, "GEMM: null pointer");
  if (epi == EPI_NONE && !scatter) {
    // compute/sgemm.go:80-96 small-matrix kernels (NN contiguous only; the bTrans branch :98-111 multiplies by
    // B as stored, a reference defect that is deliberately NOT reproduced -- see DESIGN.md)
    if (!ta && !tb && lda == k && ldb == n && ldc == n && m == n && n == k && (m == 4 || m == 8 || m == 16)) {
      if (m == 4) sgemm_small_kernel<4><<<1, 16, 0, s>>>(alpha, a, b, beta, c);
      else if (m == 8) sgemm_small_kernel<8><<<1, 64, 0, s>>>(alpha, a, b, beta, c);
      else sgemm_small_kernel<16><<<1, 256, 0, s>>>(alpha, a, b, beta, c);
      return check_launch("sgemm_small");
    }
    if ((alpha == 0.0f || k == 0) && beta == 1.0f) return GUDA_OK;        // compute/sgemm.go:139
    if (!force_exact && sgemm_mode() == GUDA_SGEMM_3XTF32) {
      rc = sgemm_3xtf32_try(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, s);
      if (rc >= 0) return rc;
    }
  }
  if (scatter && (ta || tb)) return set_error(GUDA_ERR_NOT_IMPLEMENTED, "sgemm_scatter: only the NN form is implemented");
  if (!tb && !(alpha == 0.0f && beta == 1.0f)) {
    rc = sgemm_pipe_try(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, epi, scatter, s);
    if (rc >= 0) return rc;
  }
  GemmParams<float> p{m, n, k, alpha, beta, a, lda, b, ldb, c, ldc, bias, m, n};
  if (scatter) {
    ScatterParams<float, true> sp;
    sp.n = scatter->n;
    for (int i = 0; i < GUDA_MAX_SCATTER; i++) sp.dst[i] = i < scatter->n ? scatter->dst[i] : nullptr;
    sp.ld = scatter->ld; sp.off = scatter->off;
    return launch_simt<float, false, false, MODE_FMA, EPI_NONE, true>(p, s, sp);
  }
  if (epi == EPI_BIAS_RELU) return launch_simt<float, false, false, MODE_FMA, EPI_BIAS_RELU>(p, s);
  if (epi == EPI_BIAS_GELU) return launch_simt<float, false, false, MODE_FMA, EPI_BIAS_GELU>(p, s);
  if (!ta && !tb) return launch_simt<float, false, false, MODE_FMA, EPI_NONE>(p, s);
  if (ta && !tb)  return launch_simt<float, true, false, MODE_FMA, EPI_NONE>(p, s);
  if (!ta && tb)  return launch_simt<float, false, true, MODE_FMA, EPI_NONE>(p, s);     // DotUnitary form: tolerance, not bit-exact
  return launch_simt<float, true, true, MODE_UNFUSED, EPI_NONE>(p, s);                  // AxpyInc: unfused chain
}

static int dgemm_impl(int ta, int tb, int m, int n, int k, double alpha, const double* a, int lda, const double* b, int ldb,
                      double beta, double* c, int ldc, guda_stream_t stream, const ScatterArgsT<double>* scatter = nullptr) {
  int rc = check_gemm_args(ta, tb, m, n, k, lda, ldb, ldc);
  if (rc) return rc;
  if (m == 0 || n == 0) return GUDA_OK;
  rc = bind_device();
  if (rc) return rc;
  if ((!c && !(scatter && beta == 0.0)) || ((!a || !b) && k > 0)) return set_error(GUDA_ERR_INVALID_ARG, "Dgemm: null pointer");
  cudaStream_t s = native(stream);
  GemmParams<double> p{m, n, k, alpha, beta, a, lda, b, ldb, c, ldc, nullptr, m, n};
  if (scatter) {
    // fused all-gather epilogue (sharded.cu): NN form on blocks whose edges need no separate pass
    if (ta || tb) return set_error(GUDA_ERR_NOT_IMPLEMENTED, "dgemm_scatter: only the NN form is implemented");
    if (m % 4 || n % 4) return set_error(GUDA_ERR_NOT_IMPLEMENTED, "dgemm_scatter: m and n must be multiples of 4");
    const bool rows_al = (!c || ((uintptr_t)c % 32 == 0 && (m == 1 || ((long long)ldc * 8) % 32 == 0))) &&
                         (scatter->ld * 8) % 32 == 0 && (scatter->off * 8) % 32 == 0;
    for (int i = 0; i < scatter->n; i++) if ((uintptr_t)scatter->dst[i] % 32) return set_error(GUDA_ERR_INVALID_ARG, "dgemm_scatter: destinations must be 32-byte aligned");
    if (!rows_al) return set_error(GUDA_ERR_INVALID_ARG, "dgemm_scatter: C rows must be 32-byte aligned");
    // alpha == 1: dgemm_fixed.go's plain FMA tiles (no zero test); otherwise the AXPY chain with its `tmp != 0` test
    rc = dgemm_dmma_try(0, 0, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, m, n, alpha != 1.0, s, scatter);
    if (rc >= 0) return rc;
    ScatterParams<double, true> sp;
    sp.n = scatter->n;
    for (int i = 0; i < GUDA_MAX_SCATTER; i++) sp.dst[i] = i < scatter->n ? scatter->dst[i] : nullptr;
    sp.ld = scatter->ld; sp.off = scatter->off;
    return launch_simt<double, false, false, MODE_FMA, EPI_NONE, true>(p, s, sp);        // same route guda_dgemm takes when DMMA declines
  }
  if ((alpha == 0.0 || k == 0) && beta == 1.0) return GUDA_OK;            // compute/dgemm.go:99
  // AxpyUnitary (fp64) is all-FMA only when every C row is 32-byte aligned and n % 4 == 0
  const bool rows_aligned = ((uintptr_t)c % 32 == 0) && (m == 1 || ((long long)ldc * 8) % 32 == 0) && (n % 4 == 0);
  if (!ta && !tb) {
    if (alpha == 1.0) {
      // compute/dgemm.go:243-247 -> dgemm_fixed.go: 8x8/4x4 FMA tiles + unfused edges
      int mi4 = m / 4 * 4, ni4 = n / 4 * 4;
      p.edge_m = mi4; p.edge_n = ni4;
      if (mi4 < m || ni4 < n) {
        long long edges = (long long)(m - mi4) * n + (long long)mi4 * (n - ni4);
        unsigned grid = (unsigned)((edges + 127) / 128);
        if (grid > (unsigned)sm_count() * 32) grid = sm_count() * 32;
        dgemm_edge_kernel<<<grid, 128, 0, s>>>(m, n, k, a, lda, b, ldb, beta, c, ldc, mi4, ni4);
        rc = check_launch("dgemm_edge");
        if (rc) return rc;
      }
      if (mi4 == 0 || ni4 == 0) return GUDA_OK;
      // dgemm_fixed.go's 8x8 / 4x4 kernels have no zero test: a plain FMA chain
      rc = dgemm_dmma_try(0, 0, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, mi4, ni4, false, s);
      if (rc >= 0) return rc;
      return launch_simt<double, false, false, MODE_FMA, EPI_NONE>(p, s);
    }
    if (rows_aligned) {
      rc = dgemm_dmma_try(0, 0, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, m, n, true, s);
      if (rc >= 0) return rc;
    }
    return rows_aligned ? launch_simt<double, false, false, MODE_FMA, EPI_NONE>(p, s)
                        : launch_simt<double, false, false, MODE_MIXED, EPI_NONE>(p, s);
  }
  if (ta && !tb) {
    if (rows_aligned) {
      rc = dgemm_dmma_try(1, 0, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, m, n, true, s);
      if (rc >= 0) return rc;
    }
    return rows_aligned ? launch_simt<double, true, false, MODE_FMA, EPI_NONE>(p, s)
                        : launch_simt<double, true, false, MODE_MIXED, EPI_NONE>(p, s);
  }
  if (!ta && tb) return launch_simt<double, false, true, MODE_FMA, EPI_NONE>(p, s);     // DotUnitary form: tolerance
  return launch_simt<double, true, true, MODE_UNFUSED, EPI_NONE>(p, s);
}

// fp32 GEMM with the arithmetic of bfloat16.go:136-157 (sum = beta*c; sum += (alpha*a)*b, unfused, no zero test); NN only
int sgemm_unfused_noskip(int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb, float beta,
                         float* c, int ldc, cudaStream_t s) {
  GemmParams<float> p{m, n, k, alpha, beta, a, lda, b, ldb, c, ldc, nullptr, m, n};
  return launch_simt<float, false, false, MODE_UNFUSED_NOSKIP, EPI_NONE>(p, s);
}
// exact-order SGEMM on an explicit cudaStream_t (used by the fp16 path: float16.go:231)
int sgemm_exact_on(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb, float beta,
                   float* c, int ldc, cudaStream_t s) {
  return sgemm_impl(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, nullptr, EPI_NONE, s, nullptr, true);
}

// x[i] = act(x[i] + bias[i % bias_len])   (fused_gelu.go:12-41, fused.go:220-260)
template <int EPI>
__global__ void add_bias_act_kernel(float* __restrict__ x, const float* __restrict__ bias, long long n, int bias_len) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    x[i] = apply_epilogue<float, EPI>(x[i], bias[i % bias_len]);
}
int add_bias_act_on(float* x, const float* bias, long long n, int bias_len, cudaStream_t s, bool gelu) {
  if (n < 0 || bias_len <= 0) return set_error(GUDA_ERR_INVALID_ARG, "AddBias: bad length");
  if (n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!x || !bias) return set_error(GUDA_ERR_INVALID_ARG, "AddBias: null pointer");
  long long grid = (n + 255) / 256;
  if (grid > (long long)sm_count() * 16) grid = (long long)sm_count() * 16;
  if (gelu) add_bias_act_kernel<EPI_BIAS_GELU><<<(unsigned)grid, 256, 0, s>>>(x, bias, n, bias_len);
  else add_bias_act_kernel<EPI_BIAS_RELU><<<(unsigned)grid, 256, 0, s>>>(x, bias, n, bias_len);
  return check_launch("add_bias_act");
}

}  // namespace guda

using namespace guda;
extern "C" {

int guda_sgemm(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb, float beta,
               float* c, int ldc, guda_stream_t stream) {
  return sgemm_impl(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, nullptr, EPI_NONE, native(stream));
}
int guda_dgemm(int ta, int tb, int m, int n, int k, double alpha, const double* a, int lda, const double* b, int ldb, double beta,
               double* c, int ldc, guda_stream_t stream) {
  return dgemm_impl(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream);
}

// SGEMM whose epilogue also stores the finished block at (row0, col0) of `ndest` larger row-major matrices (leading
// dimension ld_dest) -- the building block of the fused GEMM + all-gather (sharded.cu), exact-order route only.
// DGEMM twin of guda_sgemm_scatter (NN, m and n multiples of 4, 32-byte aligned rows): the DMMA kernel's epilogue stores
// the finished block at (row0, col0) of every destination matrix
int guda_dgemm_scatter(int ta, int tb, int m, int n, int k, double alpha, const double* a, int lda, const double* b, int ldb, double beta,
                       double* c, int ldc, double* const* host_dests, int ndest, int ld_dest, int row0, int col0, guda_stream_t stream) {
  if (ndest < 0 || ndest > GUDA_MAX_SCATTER || (ndest > 0 && !host_dests)) return set_error(GUDA_ERR_INVALID_ARG, "dgemm_scatter: bad destination list");
  if (ndest == 0) return dgemm_impl(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream);
  if (row0 < 0 || col0 < 0 || ld_dest < col0 + n) GUDA_PANIC("blas: bad leading dimension of C");
  ScatterArgsT<double> sc{};
  sc.n = ndest;
  for (int i = 0; i < ndest; i++) {
    if (!host_dests[i]) return set_error(GUDA_ERR_INVALID_ARG, "dgemm_scatter: null destination");
    sc.dst[i] = host_dests[i];
  }
  sc.ld = ld_dest; sc.off = (long long)row0 * ld_dest + col0;
  return dgemm_impl(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream, &sc);
}

int guda_sgemm_scatter(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb, float beta,
                       float* c, int ldc, float* const* host_dests, int ndest, int ld_dest, int row0, int col0, guda_stream_t stream) {
  if (ndest < 0 || ndest > GUDA_MAX_SCATTER || (ndest > 0 && !host_dests)) return set_error(GUDA_ERR_INVALID_ARG, "sgemm_scatter: bad destination list");
  if (ndest == 0) return sgemm_impl(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, nullptr, EPI_NONE, native(stream));
  if (row0 < 0 || col0 < 0 || ld_dest < col0 + n) GUDA_PANIC("blas: bad leading dimension of C");
  ScatterArgs sc{};
  sc.n = ndest;
  for (int i = 0; i < ndest; i++) {
    if (!host_dests[i]) return set_error(GUDA_ERR_INVALID_ARG, "sgemm_scatter: null destination");
    sc.dst[i] = host_dests[i];
  }
  sc.ld = ld_dest; sc.off = (long long)row0 * ld_dest + col0;
  return sgemm_impl(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, nullptr, EPI_NONE, native(stream), &sc);
}

static int fused_checks(int ta, int tb, float beta, const float* bias) {
  if (ta || tb) return set_error(GUDA_ERR_NOT_IMPLEMENTED, "operation not supported");   // fused.go:324-326
  if (beta != 0) return set_error(GUDA_ERR_NOT_IMPLEMENTED, "operation not supported");  // fused.go:327-329
  if (!bias) return set_error(GUDA_ERR_INVALID_ARG, "null bias");
  return GUDA_OK;
}
int guda_sgemm_bias_relu(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb,
                         float beta, float* c, int ldc, const float* bias, guda_stream_t stream) {
  int rc = fused_checks(ta, tb, beta, bias);
  if (rc) return rc;
  return sgemm_impl(0, 0, m, n, k, alpha, a, lda, b, ldb, 0.0f, c, ldc, bias, EPI_BIAS_RELU, native(stream));
}
int guda_sgemm_bias_gelu(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb,
                         float beta, float* c, int ldc, const float* bias, guda_stream_t stream) {
  int rc = fused_checks(ta, tb, beta, bias);
  if (rc) return rc;
  return sgemm_impl(0, 0, m, n, k, alpha, a, lda, b, ldb, 0.0f, c, ldc, bias, EPI_BIAS_GELU, native(stream));
}

static int add_bias_act(float* x, const float* bias, int n, int bias_len, guda_stream_t stream, bool gelu) {
  return add_bias_act_on(x, bias, n, bias_len, native(stream), gelu);
}
int guda_add_bias_gelu(float* x, const float* bias, int n, int bias_len, guda_stream_t stream) {
  return add_bias_act(x, bias, n, bias_len, stream, true);
}
int guda_add_bias_relu(float* x, const float* bias, int n, int bias_len, guda_stream_t stream) {
  return add_bias_act(x, bias, n, bias_len, stream, false);
}

}  // extern "C"
