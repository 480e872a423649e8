// dgemm_dmma.cu -- DGEMM on the FP64 tensor pipe (DMMA: mma.sync.m8n8k4.f64), NN and TN forms.
//
// Replaces compute/dgemm.go:241-260,321-333 + compute/dgemm_fixed.go:6-68 + compute/asm/f64/gemm_kernel_{4x4,8x8}_amd64.s.
// tcgen05 has no FP64 kind: on sm_100a the FP64 tensor pipe is reached through warp-level DMMA.
//
// Arithmetic contract -- identical to the SIMT kernel of gemm_simt.cuh, element by element:
//   acc = beta-scaled c;  for l ascending:  t = fl(alpha * a_il);  [if t != 0]  acc = fma(t, b_lj, acc)
// One m8n8k4 DMMA performs, for each of its 64 outputs, the four fused multiply-adds of its k-slice in ascending k
// order with one rounding each (checked bit-for-bit against the oracle's fma() chain by tests/test_gpu_dgemm_dmma.py),
// so chaining DMMAs over k reproduces the reference's sequential-k FMA chain exactly.
// The reference's zero test (`if tmp != 0`, compute/dgemm.go:256; absent on the alpha == 1 kernel route,
// dgemm_fixed.go) only changes a result when a skipped term would have met a non-finite b, or when the sum is an exact
// zero whose sign the skipped +0 would flip.  The kernel watches for exactly those cases (zero in the scaled A tile AND
// (non-finite B value OR a final accumulator equal to zero)) and then recomputes the CTA's tile with the plain
// predicated loop -- correctness insurance that costs one compare per fragment load on the fast path.
//
// Tiling: CTA 128 x 128 x 16, 256 threads = 8 warps as 2 (m) x 4 (n), warp tile 64 x 32 = 8 x 4 DMMA tiles
// (64 fp64 accumulators per thread).  Operands stream global -> shared with 16-byte cp.async (zero-fill at the edges)
// through a 4-stage ring; shared rows are padded by 4 doubles, which makes every fragment load (LDS.64, quarter-warp
// = 4 rows x 4 k) bank-conflict free.  The last K % 4 terms are applied with scalar FMAs on the accumulator fragments.
#include "gemm_simt.cuh"
#include <cstdlib>      // ScatterParams / ScatterArgsT (the fused all-gather epilogue shared with the SGEMM kernels)

namespace guda {

struct DmmaParams {
  int m, n, k;
  double alpha, beta;
  const double* a; long long lda;
  const double* b; long long ldb;
  double* c; long long ldc;
  int edge_m, edge_n;       // rows >= edge_m / cols >= edge_n are not stored (alpha == 1 NN edges belong to dgemm_edge_kernel)
  int tiles_m, tiles_n;
};

namespace dmma {
constexpr int BM = 128, BN = 128;
// Variants of the 128 x 128 CTA tile (GUDA_DMMA_WARPS=0 selects round 1's kernel as the A/B arm; evidence in DESIGN.md section 4 and
// profiles/r02/ncu_dgemm_dmma_4096_*.md).  8 warps as 2 x 4, warp tile 64 x 32 (8 x 4 DMMA tiles, 64 accumulators per thread):
//   W = 0: the next stage's cp.async issued right after the barrier (generic, clamped addresses) -- tensor pipe 81.6 % busy, 30.0
//          TFLOP/s at 8192^3: after every barrier all eight warps ran ~125 instructions of address code, then their fragment loads,
//          before the first DMMA of the tile
//   W = 2: the cp.async of the next stage is issued AFTER the first k4 step's DMMAs have been queued, interior tiles take a
//          clamp-free address path (two pointers + constant steps), and the DMMA asm is not `volatile`, so ptxas may move
//          fragment loads and address arithmetic across it -- tensor pipe 92.3 %, 33.9 TFLOP/s at 8192^3, 34.2 at 16384^3 (default)
// Measured and dropped: 16 warps of 32 x 32 (28.1: more shared-memory traffic, spills at 128 registers); W = 2 with BK 32 / 3 stages
// (32.0); W = 2 with 5 or 6 stages (33.8 / 33.7).  A register-only loop of the same DMMA pattern sustains 37.1 TFLOP/s with these
// 8 warps per SM (tools/dmma_lab.cu), so occupancy is not the limit.
template <int W> struct Warps;
template <> struct Warps<0> { static constexpr int WM = 2, WN = 4, MF = 8, NF = 4, THREADS = 256, BK = 16, STAGES = 4; static constexpr bool LATE = false; };
template <> struct Warps<2> { static constexpr int WM = 2, WN = 4, MF = 8, NF = 4, THREADS = 256, BK = 16, STAGES = 4; static constexpr bool LATE = true; };
template <int W> struct Geo {
  static constexpr int BK = Warps<W>::BK, STAGES = Warps<W>::STAGES;
  static constexpr int LDK = BK + 4;        // A (not transposed) in shared: [BM][LDK]
  static constexpr int LDT = 128 + 4;       // k-major operands in shared: [BK][LDT]
  static constexpr int A_STAGE = (BM * LDK > BK * LDT) ? BM * LDK : BK * LDT;     // doubles
  static constexpr int B_STAGE = BK * LDT;
  static constexpr int STAGE = A_STAGE + B_STAGE;
  static constexpr int SMEM_BYTES = STAGES * STAGE * 8;
};

__device__ __forceinline__ void cp_async16(double* smem, const double* g, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(g), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void dmma_8x8x4(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
// the same instruction without `volatile`: the compiler may move fragment loads and address arithmetic across it (variants 2, 3)
__device__ __forceinline__ void dmma_8x8x4_free(double (&c)[2], double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
__device__ __forceinline__ int clamp16(long long bytes) { return bytes >= 16 ? 16 : (bytes > 0 ? (int)bytes : 0); }
}  // namespace dmma

// TA: A is stored transposed ([k][m]).  SCALE: alpha != 1 (A fragments are multiplied by alpha as they are read).
// SKIP: the reference route has the `tmp != 0` test.
// SCAT: besides c (which may then be null) every finished element is stored at (row, col) + off of each matrix in sp.dst[] --
// the gathered C of every rank, peers reached through NVLink-mapped pointers (sharded.cu, guda_dgemm_sharded).
template <bool TA, bool SCALE, bool SKIP, bool SCAT = false, int W = 0>
__global__ void __launch_bounds__(dmma::Warps<W>::THREADS, 1) dgemm_dmma_kernel(DmmaParams p, ScatterParams<double, SCAT> sp) {
  using namespace dmma;
  constexpr int THREADS = Warps<W>::THREADS, MF = Warps<W>::MF, NF = Warps<W>::NF, WROWS = MF * 8, WCOLS = NF * 8;
  constexpr int BK = Geo<W>::BK, STAGES = Geo<W>::STAGES, LDK = Geo<W>::LDK, LDT = Geo<W>::LDT, A_STAGE = Geo<W>::A_STAGE, STAGE = Geo<W>::STAGE;
  constexpr bool LATE = Warps<W>::LATE;
  constexpr int CHUNKS = 64 * BK / THREADS;                 // 16-byte cp.async per thread per operand per stage
  extern __shared__ __align__(16) double dsm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp / Warps<W>::WN, wn = warp % Warps<W>::WN;
  const int g = lane >> 2, q = lane & 3;                    // fragment row group / k (or column pair) index

  // tile order: groups of 8 M-blocks sweep N so that co-resident CTAs share A rows and B columns in L2
  int tm, tn;
  {
    const int group = 8, t = blockIdx.x;
    const int per_group = group * p.tiles_n;
    const int gi = t / per_group, r = t % per_group;
    const int rows = min(group, p.tiles_m - gi * group);
    tm = gi * group + r % rows;
    tn = r / rows;
  }
  const int m0 = tm * BM, n0 = tn * BN;
  const int KT = (p.k + BK - 1) / BK;

  auto load_stage = [&](int kt, int slot) {
    double* As = dsm + slot * STAGE;
    double* Bs = As + A_STAGE;
    const int k0 = kt * BK;
#pragma unroll
    for (int it = 0; it < CHUNKS; it++) {
      const int c = tid + it * THREADS;
      if (!TA) {
        const int row = c / (BK / 2), kc = (c % (BK / 2)) * 2;
        const long long gr = m0 + row;
        const int bytes = gr < p.m ? clamp16((long long)(p.k - (k0 + kc)) * 8) : 0;
        cp_async16(As + row * LDK + kc, bytes ? p.a + gr * p.lda + (k0 + kc) : p.a, bytes);
      } else {
        const int kr = c >> 6, mc = (c & 63) * 2;
        const int bytes = (k0 + kr) < p.k ? clamp16((long long)(p.m - (m0 + mc)) * 8) : 0;
        cp_async16(As + kr * LDT + mc, bytes ? p.a + (long long)(k0 + kr) * p.lda + (m0 + mc) : p.a, bytes);
      }
    }
#pragma unroll
    for (int it = 0; it < CHUNKS; it++) {
      const int c = tid + it * THREADS;
      const int kr = c >> 6, nc = (c & 63) * 2;
      const int bytes = (k0 + kr) < p.k ? clamp16((long long)(p.n - (n0 + nc)) * 8) : 0;
      cp_async16(Bs + kr * LDT + nc, bytes ? p.b + (long long)(k0 + kr) * p.ldb + (n0 + nc) : p.b, bytes);
    }
  };

  // Interior tiles (no edge in m, n; the k edge is tested per tile): every chunk is a full 16 bytes at base + it * step -- two
  // pointers and no clamping instead of the generic path's per-chunk bounds arithmetic (~125 instructions per stage).
  const bool full_mn = LATE && m0 + BM <= p.m && n0 + BN <= p.n;
  const double* a_thr = TA ? p.a + (long long)(tid >> 6) * p.lda + (m0 + (tid & 63) * 2)
                           : p.a + (long long)(m0 + tid / (BK / 2)) * p.lda + (tid % (BK / 2)) * 2;
  const double* b_thr = p.b + (long long)(tid >> 6) * p.ldb + (n0 + (tid & 63) * 2);
  const int a_dst = TA ? (tid >> 6) * LDT + (tid & 63) * 2 : (tid / (BK / 2)) * LDK + (tid % (BK / 2)) * 2;
  const int b_dst = (tid >> 6) * LDT + (tid & 63) * 2;
  auto load_stage_fast = [&](int kt, int slot) {
    double* As = dsm + slot * STAGE;
    double* Bs = As + A_STAGE;
    const long long k0 = (long long)kt * BK;
#pragma unroll
    for (int it = 0; it < CHUNKS; it++) {
      if (!TA) cp_async16(As + a_dst + it * (THREADS / (BK / 2)) * LDK, a_thr + (long long)it * (THREADS / (BK / 2)) * p.lda + k0, 16);
      else cp_async16(As + a_dst + it * (THREADS / 64) * LDT, a_thr + (k0 + it * (THREADS / 64)) * p.lda, 16);
    }
#pragma unroll
    for (int it = 0; it < CHUNKS; it++) cp_async16(Bs + b_dst + it * (THREADS / 64) * LDT, b_thr + (k0 + it * (THREADS / 64)) * p.ldb, 16);
  };
  auto load_any = [&](int kt, int slot) {
    if (full_mn && (kt + 1) * BK <= p.k) load_stage_fast(kt, slot);
    else load_stage(kt, slot);
  };

  // ---- accumulators start from the beta-scaled C (the reference's separate pass, compute/dgemm.go:104-119) ----
  double acc[MF][NF][2];
#pragma unroll
  for (int mf = 0; mf < MF; mf++) {
    const int row = m0 + wm * WROWS + mf * 8 + g;
#pragma unroll
    for (int nf = 0; nf < NF; nf++) {
      const int col = n0 + wn * WCOLS + nf * 8 + q * 2;
#pragma unroll
      for (int e = 0; e < 2; e++) {
        double v = 0.0;
        if (p.beta != 0.0 && p.c && row < p.m && col + e < p.n) {
          v = p.c[(long long)row * p.ldc + col + e];
          if (p.beta != 1.0) v = __dmul_rn(v, p.beta);
        }
        acc[mf][nf][e] = v;
      }
    }
  }

#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < KT) load_stage(s, s);
    cp_async_commit();
  }

  bool zero_a = false, nonfinite_b = false;
  for (int kt = 0; kt < KT; kt++) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    if (!LATE) {
      if (kt + STAGES - 1 < KT) load_stage(kt + STAGES - 1, (kt + STAGES - 1) % STAGES);
      cp_async_commit();
    }

    const double* As = dsm + (kt % STAGES) * STAGE;
    const double* Bs = As + A_STAGE;
    const int k0 = kt * BK;
#pragma unroll
    for (int kk = 0; kk < BK / 4; kk++) {
      const int kb = k0 + kk * 4;
      if (LATE && kk == 1) {          // the slot refilled here was read in the previous iteration; every warp is past this iteration's barrier
        if (kt + STAGES - 1 < KT) load_any(kt + STAGES - 1, (kt + STAGES - 1) % STAGES);
        cp_async_commit();
      }
      if (kb + 4 <= p.k) {
        double af[MF], bf[NF];
#pragma unroll
        for (int mf = 0; mf < MF; mf++) {
          const int ml = wm * WROWS + mf * 8 + g;
          double v = TA ? As[(kk * 4 + q) * LDT + ml] : As[ml * LDK + kk * 4 + q];
          if (SCALE) v = __dmul_rn(p.alpha, v);
          if (SKIP) zero_a |= (v == 0.0);
          af[mf] = v;
        }
#pragma unroll
        for (int nf = 0; nf < NF; nf++) {
          const double v = Bs[(kk * 4 + q) * LDT + wn * WCOLS + nf * 8 + g];
          if (SKIP) nonfinite_b |= ((__double2hiint(v) & 0x7ff00000) == 0x7ff00000);
          bf[nf] = v;
        }
#pragma unroll
        for (int mf = 0; mf < MF; mf++)
#pragma unroll
          for (int nf = 0; nf < NF; nf++) {
            if (LATE) dmma_8x8x4_free(acc[mf][nf], af[mf], bf[nf]);
            else dmma_8x8x4(acc[mf][nf], af[mf], bf[nf]);
          }
      } else if (kb < p.k) {
        // K % 4 tail: scalar FMAs on the accumulator fragments, same order and the same zero test
        for (int l = kb; l < p.k; l++) {
          const int kl = l - k0;
          double av[MF];
#pragma unroll
          for (int mf = 0; mf < MF; mf++) {
            const int ml = wm * WROWS + mf * 8 + g;
            double v = TA ? As[kl * LDT + ml] : As[ml * LDK + kl];
            if (SCALE) v = __dmul_rn(p.alpha, v);
            av[mf] = v;
          }
#pragma unroll
          for (int nf = 0; nf < NF; nf++) {
#pragma unroll
            for (int e = 0; e < 2; e++) {
              const double bv = Bs[kl * LDT + wn * WCOLS + nf * 8 + q * 2 + e];
#pragma unroll
              for (int mf = 0; mf < MF; mf++)
                if (!SKIP || av[mf] != 0.0) acc[mf][nf][e] = __fma_rn(av[mf], bv, acc[mf][nf][e]);
            }
          }
        }
      }
    }
  }
  cp_async_wait<0>();

  // ---- zero-test insurance (see header) ----
  if (SKIP) {
    bool zero_out = false;
#pragma unroll
    for (int mf = 0; mf < MF; mf++)
#pragma unroll
      for (int nf = 0; nf < NF; nf++) zero_out |= (acc[mf][nf][0] == 0.0) | (acc[mf][nf][1] == 0.0);
    const int za = __syncthreads_or(zero_a);
    const int hz = __syncthreads_or(nonfinite_b | zero_out);
    if (za && hz) {
      // exact predicated loop for this CTA's tile; a warp walks 32 consecutive columns of one row
      for (int e = tid; e < BM * BN; e += THREADS) {
        const int row = m0 + e / BN, col = n0 + e % BN;
        if (row >= p.m || col >= p.n || row >= p.edge_m || col >= p.edge_n) continue;
        double v = 0.0;
        if (p.beta != 0.0 && p.c) {
          v = p.c[(long long)row * p.ldc + col];
          if (p.beta != 1.0) v = __dmul_rn(v, p.beta);
        }
        for (int l = 0; l < p.k; l++) {
          const double t = __dmul_rn(p.alpha, TA ? p.a[(long long)l * p.lda + row] : p.a[(long long)row * p.lda + l]);
          if (t != 0.0) v = __fma_rn(t, p.b[(long long)l * p.ldb + col], v);
        }
        if (!SCAT || p.c) p.c[(long long)row * p.ldc + col] = v;
        if constexpr (SCAT) {
          for (int d = 0; d < sp.n; d++) sp.dst[d][(long long)row * sp.ld + sp.off + col] = v;
        }
      }
      return;
    }
  }

  // ---- single store of C ----
  const bool vec_ok = ((p.ldc & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.c) & 15) == 0);
  const int lim_m = p.m < p.edge_m ? p.m : p.edge_m, lim_n = p.n < p.edge_n ? p.n : p.edge_n;
#pragma unroll
  for (int mf = 0; mf < MF; mf++) {
    const int row = m0 + wm * WROWS + mf * 8 + g;
    if (row >= lim_m) continue;
#pragma unroll
    for (int nf = 0; nf < NF; nf++) {
      const int col = n0 + wn * WCOLS + nf * 8 + q * 2;
      if (!SCAT || p.c) {
        double* dst = p.c + (long long)row * p.ldc + col;
        if (vec_ok && col + 1 < lim_n) {
          *reinterpret_cast<double2*>(dst) = make_double2(acc[mf][nf][0], acc[mf][nf][1]);
        } else {
          if (col < lim_n) dst[0] = acc[mf][nf][0];
          if (col + 1 < lim_n) dst[1] = acc[mf][nf][1];
        }
      }
      if constexpr (SCAT) {
        const bool svec = ((sp.ld & 1) == 0) && ((sp.off & 1) == 0) && col + 1 < lim_n;
        for (int d = 0; d < sp.n; d++) {
          double* dst = sp.dst[d] + (long long)row * sp.ld + sp.off + col;
          if (svec && (reinterpret_cast<uintptr_t>(sp.dst[d]) & 15) == 0) {
            *reinterpret_cast<double2*>(dst) = make_double2(acc[mf][nf][0], acc[mf][nf][1]);
          } else {
            if (col < lim_n) dst[0] = acc[mf][nf][0];
            if (col + 1 < lim_n) dst[1] = acc[mf][nf][1];
          }
        }
      }
    }
  }
}

static int g_dmma_warps = [] { const char* e = getenv("GUDA_DMMA_WARPS"); return e ? atoi(e) : 2; }();
template <bool TA, bool SCALE, bool SKIP, bool SCAT, int W>
static int launch_dmma_w(const DmmaParams& p, const ScatterParams<double, SCAT>& sp, cudaStream_t s) {
  auto kern = dgemm_dmma_kernel<TA, SCALE, SKIP, SCAT, W>;
  static bool attr_set = false;
  if (!attr_set) {
    GUDA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, dmma::Geo<W>::SMEM_BYTES));
    attr_set = true;
  }
  kern<<<p.tiles_m * p.tiles_n, dmma::Warps<W>::THREADS, dmma::Geo<W>::SMEM_BYTES, s>>>(p, sp);
  return check_launch("dgemm_dmma");
}
template <bool TA, bool SCALE, bool SKIP, bool SCAT>
static int launch_dmma_s(const DmmaParams& p, const ScatterParams<double, SCAT>& sp, cudaStream_t s) {
  switch (g_dmma_warps) {
    case 0: return launch_dmma_w<TA, SCALE, SKIP, SCAT, 0>(p, sp, s);
    default: return launch_dmma_w<TA, SCALE, SKIP, SCAT, 2>(p, sp, s);
  }
}
template <bool TA, bool SCALE, bool SKIP>
static int launch_dmma(const DmmaParams& p, cudaStream_t s, const ScatterArgsT<double>* sc) {
  if (!sc) return launch_dmma_s<TA, SCALE, SKIP, false>(p, {}, s);
  ScatterParams<double, true> sp;
  sp.n = sc->n;
  for (int i = 0; i < GUDA_MAX_SCATTER; i++) sp.dst[i] = i < sc->n ? sc->dst[i] : nullptr;
  sp.ld = sc->ld; sp.off = sc->off;
  return launch_dmma_s<TA, SCALE, SKIP, true>(p, sp, s);
}

static int g_dgemm_mode = GUDA_DGEMM_TENSOR;
int dgemm_dmma_one(int ta, int m, int n, int k, double alpha, const double* a, int lda, const double* b, int ldb, double beta, double* c,
                   int ldc, int edge_m, int edge_n, bool skip, cudaStream_t s, const ScatterArgsT<double>* sc);
int dgemm_mode() { return g_dgemm_mode; }

// Returns GUDA_OK / an error if the DMMA route took the call, -1 if the operands are not eligible (caller uses the SIMT kernel).
// Eligible: B not transposed, 16-byte aligned A and B with even leading dimensions (cp.async granule).
int dgemm_dmma_try(int ta, int tb, int m, int n, int k, double alpha, const double* a, int lda, const double* b, int ldb, double beta,
                   double* c, int ldc, int edge_m, int edge_n, bool skip, cudaStream_t s, const ScatterArgsT<double>* sc) {
  if (g_dgemm_mode != GUDA_DGEMM_TENSOR || tb) return -1;
  if (m < 32 || n < 32 || k < 4) return -1;
  if (((uintptr_t)a & 15) || ((uintptr_t)b & 15) || (lda & 1) || (ldb & 1)) return -1;
  if ((long long)((m + dmma::BM - 1) / dmma::BM) * ((n + dmma::BN - 1) / dmma::BN) > 0x7fffffffLL) return -1;
  // K blocking across launches keeps the co-resident CTAs' panels inside L2 (same bits: see launch_simt in gemm.cu)
  constexpr int KC = 4096;
  if (k >= 2 * KC && (long long)m * n >= (1ll << 22)) {
    for (int k0 = 0; k0 < k; k0 += KC) {
      const int kc = (k0 + KC >= k) ? k - k0 : KC;
      if (sc && !c) return set_error(GUDA_ERR_INVALID_ARG, "dgemm_scatter: a K-chunked call needs a local C block");
      int rc = dgemm_dmma_one(ta, m, n, kc, alpha, ta ? a + (long long)k0 * lda : a + k0, lda, b + (long long)k0 * ldb, ldb,
                              k0 == 0 ? beta : 1.0, c, ldc, edge_m, edge_n, skip, s, (k0 + KC >= k) ? sc : nullptr);
      if (rc) return rc;
    }
    return GUDA_OK;
  }
  return dgemm_dmma_one(ta, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, edge_m, edge_n, skip, s, sc);
}

int dgemm_dmma_one(int ta, int m, int n, int k, double alpha, const double* a, int lda, const double* b, int ldb, double beta, double* c,
                   int ldc, int edge_m, int edge_n, bool skip, cudaStream_t s, const ScatterArgsT<double>* sc) {
  DmmaParams p{m, n, k, alpha, beta, a, lda, b, ldb, c, ldc, edge_m, edge_n, (m + dmma::BM - 1) / dmma::BM, (n + dmma::BN - 1) / dmma::BN};
  const bool scale = alpha != 1.0;
  if (ta) {
    if (scale) return skip ? launch_dmma<true, true, true>(p, s, sc) : launch_dmma<true, true, false>(p, s, sc);
    return skip ? launch_dmma<true, false, true>(p, s, sc) : launch_dmma<true, false, false>(p, s, sc);
  }
  if (scale) return skip ? launch_dmma<false, true, true>(p, s, sc) : launch_dmma<false, true, false>(p, s, sc);
  return skip ? launch_dmma<false, false, true>(p, s, sc) : launch_dmma<false, false, false>(p, s, sc);
}

}  // namespace guda

extern "C" {
int guda_set_dgemm_mode(int mode) {
  if (mode != GUDA_DGEMM_TENSOR && mode != GUDA_DGEMM_SIMT) return guda::set_error(GUDA_ERR_INVALID_ARG, "unknown dgemm mode %d", mode);
  guda::g_dgemm_mode = mode;
  return GUDA_OK;
}
int guda_get_dgemm_mode(int* mode) {
  if (!mode) return guda::set_error(GUDA_ERR_INVALID_ARG, "null pointer");
  *mode = guda::g_dgemm_mode;
  return GUDA_OK;
}
}
