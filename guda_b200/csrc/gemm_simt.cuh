// gemm_simt.cuh -- exact-order SIMT GEMM for fp32 / fp64 (row-major, all four transpose forms).
//
// Arithmetic contract (matches compute/sgemm.go:139-350 and compute/dgemm.go:98-362 element by element):
//   acc   = (beta == 0) ? 0 : (beta == 1 ? c : c*beta)               the reference's separate beta pass
//   for l = 0 .. k-1 (ascending):  t = fl(alpha * a_il);  if (t != 0)  acc = CHAIN(t, b_lj, acc)
//   CHAIN = fma            (AxpyUnitary FMA route: NN, TN)                                   MODE_FMA
//         = fl(t*b) + acc  (AxpyInc, unfused: TT)                                            MODE_UNFUSED
//         = per-element mix (fp64 AxpyUnitary: alignment peel and the last n%4 are unfused)  MODE_MIXED
// Each C element is owned by ONE thread that walks k in ascending order, so the rounding sequence
// is the reference's.  The zero test is hoisted: A is scaled by alpha while it is staged into
// shared memory, and a block-wide OR (free: it rides the tile barrier) tells whether the staged
// A tile holds any zero; only then does the k-loop take the predicated path.
//
// Tiling: BM x BN x BK per CTA of 256 threads (16 x 16), each thread a (BM/16) x (BN/16) micro-tile
// split in two halves BM/2 (BN/2) apart so that shared-memory reads are 128-bit and conflict-free;
// A is stored k-major in shared memory; global->register->shared double buffering, one barrier per
// k-tile.  Epilogues (bias+ReLU / bias+GELU) are applied to the accumulators in registers before
// the single store of C (subsystem 4 of the north star; replaces the second pass of fused.go:357-367
// and fused_gelu.go:107-125).
#pragma once
#include "common.cuh"
#include <type_traits>

namespace guda {

enum { MODE_FMA = 0, MODE_UNFUSED = 1, MODE_MIXED = 2, MODE_UNFUSED_NOSKIP = 3 };   // NOSKIP: bfloat16.go:152 has no zero test
enum { EPI_NONE = 0, EPI_BIAS_RELU = 1, EPI_BIAS_GELU = 2 };
#define GUDA_MAX_SCATTER 8

template <typename T> struct GemmVec;
template <> struct GemmVec<float> { using V = float4; static constexpr int N = 4; };
template <> struct GemmVec<double> { using V = double2; static constexpr int N = 2; };

__device__ __forceinline__ float g_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double g_fma(double a, double b, double c) { return __fma_rn(a, b, c); }
__device__ __forceinline__ float g_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double g_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float g_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double g_add(double a, double b) { return __dadd_rn(a, b); }

// Two independent IEEE round-to-nearest fp32 FMAs in one instruction (fma.rn.f32x2 -> FFMA2 on sm_100a): the same bits as
// two scalar fmaf() calls, half the issue slots.  c0 = fma(a, b0, c0); c1 = fma(a, b1, c1).  ptxas keeps the accumulator
// pair and the B pair (straight out of LDS.128) in aligned 64-bit registers, so the mov.b64 packs below cost nothing.
__device__ __forceinline__ void fma2_f32(float& c0, float& c1, unsigned long long aa, float b0, float b1) {
  unsigned long long c, bb;
  asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(c0), "f"(c1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(bb) : "f"(b0), "f"(b1));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(c) : "l"(aa), "l"(bb));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(c0), "=f"(c1) : "l"(c));
}
__device__ __forceinline__ unsigned long long dup_f32(float a) {
  unsigned long long aa;
  asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
  return aa;
}

// ---- geluFloat32 restated with explicit roundings (fused_gelu.go:44-56, activations.go:31-90) ----
__device__ __forceinline__ float gelu_exp_f32(float x) {
  if (x > 88.7f) return 3.40282346638528859811704183484516925440e+38f;
  if (x < -87.3f) return 0.0f;
  const double ln2 = 0.6931471805599453094;
  // k = int(math.Floor(float64(x) / ln2)) (activations.go:71).  The fp64 division is replaced by a multiplication with the
  // correctly rounded reciprocal: an exhaustive host-side check (expcheck.c in the test infrastructure, run by the CPU test
  // suite) walks all 2,237,661,186 float32 inputs that reach this line and finds floor(x / ln2) == floor(x * (1 / ln2)) for each.
  int k = (int)floor(__dmul_rn((double)x, 0x1.71547652b82fep+0));
  float r = __fsub_rn(x, __fmul_rn((float)k, (float)ln2));
  float r2 = __fmul_rn(r, r), r3 = __fmul_rn(r2, r), r4 = __fmul_rn(r2, r2), r5 = __fmul_rn(r4, r);
  float e = __fadd_rn(1.0f, r);
  e = __fadd_rn(e, __fmul_rn(0.4999999701976776f, r2));
  e = __fadd_rn(e, __fmul_rn(0.1666666567325592f, r3));
  e = __fadd_rn(e, __fmul_rn(0.0416666679084301f, r4));
  e = __fadd_rn(e, __fmul_rn(0.0083333337679505f, r5));
  // float32(math.Ldexp(float64(e), k)) (activations.go:88): k is in [-126, 127] and e in [1, 2], so 2^k is a normal float and
  // the product rounds once from the same exact value (overflow to +Inf included) -- also covered by expcheck.c
  return __fmul_rn(e, __int_as_float((k + 127) << 23));
}
// ---- IEEE divisions of the activation functions without div.rn's generic sequence (~10 instructions + a slow-path call each) ----
// The three divisions of geluFloat32 / the one of SigmoidFloat32 made these kernels issue-bound at 0.47 of HBM.  Each activation is a
// function of ONE float, so a cheaper sequence is certified the brute-force way: guda_selftest_activations() evaluates the sequences
// below and the plain __fdiv_rn restatement on all 2^32 inputs ON THE DEVICE and counts differing bit patterns (a -m gpu test
// requires 0).  FAST = false selects the __fdiv_rn form (the other arm of that check).
// q = RN(a / D), D a small integer constant: Markstein's correction with the correctly rounded reciprocal.
template <int D> __device__ __forceinline__ float div_const_rn(float a) {
  constexpr float d = (float)D, c = 1.0f / (float)D;
  const float q0 = __fmul_rn(a, c);
  const float r = __fmaf_rn(-d, q0, a);                 // exact remainder
  return __fmaf_rn(r, c, q0);
}
// q = RN(a / b) for normal b well inside the exponent range and a / b normal: approximate reciprocal, one Newton step, quotient,
// exact remainder, correction.
__device__ __forceinline__ float div_normal_rn(float a, float b) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(b));
  const float e = __fmaf_rn(-b, y, 1.0f);
  y = __fmaf_rn(y, e, y);
  const float q0 = __fmul_rn(a, y);
  const float r = __fmaf_rn(-b, q0, a);
  return __fmaf_rn(r, y, q0);
}
template <bool FAST = true>
__device__ __forceinline__ float gelu_tanh_f32(float x) {
  if (x != x) return x;
  float ax = fabsf(x);
  float r;
  if (ax > 10.0f) r = 1.0f;
  else if (ax < 0.5f) {
    float x2 = __fmul_rn(ax, ax);
    float b = __fsub_rn(1.0f, FAST ? div_const_rn<3>(x2) : __fdiv_rn(x2, 3.0f));
    float n15 = __fmul_rn(__fmul_rn(2.0f, x2), x2);
    float c = FAST ? div_const_rn<15>(n15) : __fdiv_rn(n15, 15.0f);
    r = __fmul_rn(ax, __fadd_rn(b, c));
  } else {
    float e = gelu_exp_f32(__fmul_rn(2.0f, ax));
    r = FAST ? div_normal_rn(__fsub_rn(e, 1.0f), __fadd_rn(e, 1.0f)) : __fdiv_rn(__fsub_rn(e, 1.0f), __fadd_rn(e, 1.0f));
  }
  return x < 0 ? -r : r;
}
template <bool FAST = true>
__device__ __forceinline__ float gelu_f32_exact(float x) {
  float x3 = __fmul_rn(__fmul_rn(x, x), x);
  float inner = __fadd_rn(x, __fmul_rn(0.044715f, x3));
  float arg = __fmul_rn(0.7978845608028653559f, inner);
  float t = gelu_tanh_f32<FAST>(arg);
  return __fmul_rn(__fmul_rn(0.5f, x), __fadd_rn(1.0f, t));
}

template <typename T, int EPI>
__device__ __forceinline__ T apply_epilogue(T acc, T bias) {
  if (EPI == EPI_NONE) return acc;
  T v = g_add(acc, bias);
  if (EPI == EPI_BIAS_RELU) return v < T(0) ? T(0) : v;          // fused.go:362-365 (NaN stays NaN)
  return (T)gelu_f32_exact<true>((float)v);
}

template <typename T>
struct GemmParams {
  int m, n, k;
  T alpha, beta;
  const T* a; long long lda;
  const T* b; long long ldb;
  T* c; long long ldc;
  const T* bias;
  int edge_m, edge_n;      // elements with i >= edge_m or j >= edge_n are NOT stored (fp64 NN alpha==1 edges)
};

// fused all-gather (sharded.cu): besides c, the finished tile is stored at (row, col) + off of every matrix in dst[] -- the
// gathered C of each rank, peers reached through NVLink-mapped (CUDA IPC) pointers; c may then be null.  A separate kernel
// parameter (empty for the ordinary kernels) so that the hot kernels' parameter block and code are untouched: ptxas'
// schedule of the main loop proved sensitive to it (57.4 -> 54.6 TFLOP/s at 8192^3 when these fields lived in GemmParams).
template <typename T, bool SCAT> struct ScatterParams {};
template <typename T> struct ScatterParams<T, true> {
  int n;
  T* dst[GUDA_MAX_SCATTER];
  long long ld, off;
};

template <typename T> struct ScatterArgsT { int n; T* dst[GUDA_MAX_SCATTER]; long long ld, off; };
using ScatterArgs = ScatterArgsT<float>;

template <typename T, int BM, int BN, int BK>
struct GemmSmem {
  static constexpr int VN = GemmVec<T>::N;
  static constexpr int LDA_S = BM + VN;
  static constexpr int LDB_S = BN + VN;
  T a[2][BK][LDA_S];
  T b[2][BK][LDB_S];
};

// global -> registers for one operand tile.  KCONTIG: the k index is the contiguous one in global
// memory (A not transposed / B transposed); otherwise the m (n) index is contiguous.
template <typename T, int BMN, int BK, bool KCONTIG, bool FAST>
struct TileLoader {
  using V = typename GemmVec<T>::V;
  static constexpr int VN = GemmVec<T>::N;
  static constexpr int NVEC = BMN * BK / VN;
  static constexpr int PER_THREAD = NVEC / 256;
  static_assert(NVEC % 256 == 0, "tile must split evenly over 256 threads");
  V reg[PER_THREAD];

  // scale / do_scale: on the edge path (FAST == false) operand A is scaled HERE, element by element, so that the zero padding
  // beyond (MN, K) stays an exact +0 -- scaling it in store() would turn it into NaN for alpha = +-Inf (0 * Inf) and the
  // zero test `t != 0` would then let the NaN into every accumulator of the tile.
  __device__ __forceinline__ void load(const T* __restrict__ g, long long ld, int mn0, int k0, int MN, int K, int tid, T scale = T(1),
                                       bool do_scale = false) {
#pragma unroll
    for (int it = 0; it < PER_THREAD; it++) {
      int v = tid + it * 256;
      int mn, kk;
      if (KCONTIG) { mn = v / (BK / VN); kk = (v % (BK / VN)) * VN; }
      else         { kk = v / (BMN / VN); mn = (v % (BMN / VN)) * VN; }
      T* r = reinterpret_cast<T*>(&reg[it]);
      if (FAST) {
        const T* p = KCONTIG ? g + (long long)(mn0 + mn) * ld + (k0 + kk) : g + (long long)(k0 + kk) * ld + (mn0 + mn);
        reg[it] = *reinterpret_cast<const V*>(p);
      } else {
#pragma unroll
        for (int j = 0; j < VN; j++) {
          int gm = mn0 + mn + (KCONTIG ? 0 : j), gk = k0 + kk + (KCONTIG ? j : 0);
          bool ok = gm < MN && gk < K;
          const T* p = KCONTIG ? g + (long long)gm * ld + gk : g + (long long)gk * ld + gm;
          r[j] = ok ? (do_scale ? g_mul(scale, *p) : *p) : T(0);
        }
      }
    }
  }
  // registers -> shared (k-major [BK][LD]); returns true if any stored value is zero (after scaling)
  template <int LD>
  __device__ __forceinline__ bool store(T (*s)[LD], T scale, bool do_scale, int tid) {
    bool anyzero = false;
#pragma unroll
    for (int it = 0; it < PER_THREAD; it++) {
      int v = tid + it * 256;
      int mn, kk;
      if (KCONTIG) { mn = v / (BK / VN); kk = (v % (BK / VN)) * VN; }
      else         { kk = v / (BMN / VN); mn = (v % (BMN / VN)) * VN; }
      T* r = reinterpret_cast<T*>(&reg[it]);
      if (do_scale) {
#pragma unroll
        for (int j = 0; j < VN; j++) { if (FAST) r[j] = g_mul(scale, r[j]); anyzero |= (r[j] == T(0)); }
      }
      if (KCONTIG) {
#pragma unroll
        for (int j = 0; j < VN; j++) s[kk + j][mn] = r[j];
      } else {
        *reinterpret_cast<V*>(&s[kk][mn]) = reg[it];
      }
    }
    return anyzero;
  }
};

template <typename T, int H>
__device__ __forceinline__ void load_frag(T* dst, const T* src) {
  constexpr int BYTES = H * (int)sizeof(T);
  if constexpr (BYTES % 16 == 0) {
#pragma unroll
    for (int i = 0; i < BYTES / 16; i++) reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
  } else if constexpr (BYTES == 8) {
    *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(src);
  } else {
#pragma unroll
    for (int i = 0; i < H; i++) dst[i] = src[i];
  }
}

// TXN threads along n (256 / TXN along m): 16 -> the square grid; 32 -> a warp is one m-row of threads, every A fragment load is
// a warp-wide broadcast and the B fragments are 8-byte loads (micro-tile 8 x 4 on the 64 x 128 tile).
template <typename T, bool TA, bool TB, int BM, int BN, int BK, int MODE, int EPI, bool FAST, int MINB, bool SCAT = false, int TXN = 16>
__global__ void __launch_bounds__(256, MINB) gemm_simt_kernel(GemmParams<T> p, ScatterParams<T, SCAT> sp) {
  constexpr int TM = BM / (256 / TXN), TN = BN / TXN;       // micro-tile
  constexpr int HM = TM / 2, HN = TN / 2;         // halves, BM/2 (BN/2) apart
  using Smem = GemmSmem<T, BM, BN, BK>;
  __shared__ __align__(16) Smem sm;

  const int tid = threadIdx.x;
  const int tx = tid % TXN, ty = tid / TXN;
  // CTA -> tile: groups of 16 M-blocks sweep N, so the CTAs resident at one time cover a ~16 x 18 patch of C and
  // share A row panels and B column panels in L2 (row-major CTA order re-read B once per 4-5 tile rows:
  // 8.1 GB of DRAM reads for 8192^3 against 0.8 GB algorithmic; profiles/r01)
  int tile_m, tile_n;
  {
    const int tiles_n = gridDim.x, tiles_m = gridDim.y, group = 16;
    const int t = blockIdx.y * tiles_n + blockIdx.x;
    const int per_group = group * tiles_n;
    const int gi = t / per_group, r = t - gi * per_group;
    const int rows = min(group, tiles_m - gi * group);
    tile_m = gi * group + r % rows;
    tile_n = r / rows;
  }
  const int m0 = tile_m * BM, n0 = tile_n * BN;

  TileLoader<T, BM, BK, !TA, FAST> la;
  TileLoader<T, BN, BK, TB, FAST> lb;

  // ---- accumulators start from the beta-scaled C (the reference's separate pass over C) ----
  T acc[TM][TN];
  unsigned long long fused_mask = ~0ull;          // MODE_MIXED: bit (i*TN+j) set -> FMA
#pragma unroll
  for (int i = 0; i < TM; i++) {
    int row = m0 + (i < HM ? ty * HM + i : BM / 2 + ty * HM + (i - HM));
    long long peel = 0, fhi = 0;
    if (MODE == MODE_MIXED) {
      // f64/axpyunitary_fma_amd64.s: scalar (unfused) until &c[row][j] is 32-byte aligned, FMA for the
      // 16- and 4-wide bodies, unfused for the last (n-peel)%4 elements
      unsigned long long addr = (unsigned long long)(p.c + (long long)row * p.ldc);
      peel = ((32 - (addr & 31)) & 31) / sizeof(T);
      if (peel > p.n) peel = p.n;
      fhi = peel + ((p.n - peel) & ~3ll);
    }
#pragma unroll
    for (int j = 0; j < TN; j++) {
      int col = n0 + (j < HN ? tx * HN + j : BN / 2 + tx * HN + (j - HN));
      T v = T(0);
      if (EPI == EPI_NONE && p.beta != T(0) && row < p.m && col < p.n) {
        v = p.c[(long long)row * p.ldc + col];
        if (p.beta != T(1)) v = g_mul(v, p.beta);
      }
      acc[i][j] = v;
      if (MODE == MODE_MIXED && !(col >= peel && col < fhi)) fused_mask &= ~(1ull << (i * TN + j));
    }
  }

  const int KT = (p.k + BK - 1) / BK;
  bool zf = false;
  if (KT > 0) {
    la.load(p.a, p.lda, m0, 0, p.m, p.k, tid, p.alpha, true);
    lb.load(p.b, p.ldb, n0, 0, p.n, p.k, tid);
    bool z = la.template store<Smem::LDA_S>(sm.a[0], p.alpha, true, tid);
    lb.template store<Smem::LDB_S>(sm.b[0], T(0), false, tid);
    zf = __syncthreads_or(z);
  }
  for (int kt = 0; kt < KT; kt++) {
    const int cur = kt & 1;
    if (kt + 1 < KT) {
      la.load(p.a, p.lda, m0, (kt + 1) * BK, p.m, p.k, tid, p.alpha, true);
      lb.load(p.b, p.ldb, n0, (kt + 1) * BK, p.n, p.k, tid);
    }
    if (!zf && MODE == MODE_FMA) {
#pragma unroll
      for (int kk = 0; kk < BK; kk++) {
        T af[TM], bf[TN];
        load_frag<T, HM>(af, &sm.a[cur][kk][ty * HM]);
        load_frag<T, HM>(af + HM, &sm.a[cur][kk][BM / 2 + ty * HM]);
        load_frag<T, HN>(bf, &sm.b[cur][kk][tx * HN]);
        load_frag<T, HN>(bf + HN, &sm.b[cur][kk][BN / 2 + tx * HN]);
        if constexpr (sizeof(T) == 4 && TXN == 32 && TM % 2 == 0) {
          // tall micro-tile: pair the accumulators along m, so the duplicated operand is the (few) B values and the A pairs
          // come straight out of the 128-bit fragment loads; fma(b, a, c) == fma(a, b, c)
#pragma unroll
          for (int j = 0; j < TN; j++) {
            const unsigned long long bb = dup_f32((float)bf[j]);
#pragma unroll
            for (int i = 0; i < TM; i += 2)
              fma2_f32(reinterpret_cast<float&>(acc[i][j]), reinterpret_cast<float&>(acc[i + 1][j]), bb, (float)af[i], (float)af[i + 1]);
          }
        } else if constexpr (sizeof(T) == 4 && TN % 2 == 0) {
#pragma unroll
          for (int i = 0; i < TM; i++) {
            const unsigned long long aa = dup_f32((float)af[i]);
#pragma unroll
            for (int j = 0; j < TN; j += 2)
              fma2_f32(reinterpret_cast<float&>(acc[i][j]), reinterpret_cast<float&>(acc[i][j + 1]), aa, (float)bf[j], (float)bf[j + 1]);
          }
        } else {
#pragma unroll
          for (int i = 0; i < TM; i++)
#pragma unroll
            for (int j = 0; j < TN; j++) acc[i][j] = g_fma(af[i], bf[j], acc[i][j]);
        }
      }
    } else {
      const int kreal = FAST ? BK : (p.k - kt * BK < BK ? p.k - kt * BK : BK);     // the reference never visits k >= K (bfloat16.go:152 has no zero test to hide it)
#pragma unroll
      for (int kk = 0; kk < BK; kk++) {
        if (!FAST && kk >= kreal) break;
        T af[TM], bf[TN];
        load_frag<T, HM>(af, &sm.a[cur][kk][ty * HM]);
        load_frag<T, HM>(af + HM, &sm.a[cur][kk][BM / 2 + ty * HM]);
        load_frag<T, HN>(bf, &sm.b[cur][kk][tx * HN]);
        load_frag<T, HN>(bf + HN, &sm.b[cur][kk][BN / 2 + tx * HN]);
#pragma unroll
        for (int i = 0; i < TM; i++) {
          if (MODE == MODE_UNFUSED_NOSKIP || af[i] != T(0)) {            // compute/sgemm.go:301 `if tmp != 0` (true for NaN)
#pragma unroll
            for (int j = 0; j < TN; j++) {
              if (MODE == MODE_FMA) acc[i][j] = g_fma(af[i], bf[j], acc[i][j]);
              else if (MODE == MODE_UNFUSED || MODE == MODE_UNFUSED_NOSKIP) acc[i][j] = g_add(g_mul(bf[j], af[i]), acc[i][j]);
              else acc[i][j] = ((fused_mask >> (i * TN + j)) & 1ull) ? g_fma(bf[j], af[i], acc[i][j])
                                                                    : g_add(g_mul(bf[j], af[i]), acc[i][j]);
            }
          }
        }
      }
    }
    bool z = false;
    if (kt + 1 < KT) {
      z = la.template store<Smem::LDA_S>(sm.a[cur ^ 1], p.alpha, true, tid);
      lb.template store<Smem::LDB_S>(sm.b[cur ^ 1], T(0), false, tid);
    }
    zf = __syncthreads_or(z);
  }

  // ---- epilogue in registers, single store of C ----
  using V = typename GemmVec<T>::V;
  constexpr int VN = GemmVec<T>::N;
#pragma unroll
  for (int i = 0; i < TM; i++) {
    int row = m0 + (i < HM ? ty * HM + i : BM / 2 + ty * HM + (i - HM));
    if (row >= p.m || row >= p.edge_m) continue;
#pragma unroll
    for (int jh = 0; jh < 2; jh++) {
      int col0 = n0 + (jh == 0 ? tx * HN : BN / 2 + tx * HN);
      T out[HN];
#pragma unroll
      for (int j = 0; j < HN; j++) {
        T bias = T(0);
        if (EPI != EPI_NONE) bias = (col0 + j < p.n) ? p.bias[col0 + j] : T(0);
        out[j] = apply_epilogue<T, EPI>(acc[i][jh * HN + j], bias);
      }
      int lim = p.n < p.edge_n ? p.n : p.edge_n;
      if (!SCAT || p.c) {
        T* dst = p.c + (long long)row * p.ldc + col0;
        if (FAST && HN % VN == 0) {
#pragma unroll
          for (int j = 0; j < HN; j += VN) *reinterpret_cast<V*>(dst + j) = *reinterpret_cast<V*>(out + j);
        } else if (FAST && sizeof(T) == 4 && HN == 2) {
          *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<uint2*>(out);       // 32 x 8 thread grid: a warp stores 256 contiguous bytes
        } else {
#pragma unroll
          for (int j = 0; j < HN; j++) if (col0 + j < lim) dst[j] = out[j];
        }
      }
      if constexpr (SCAT) {
        const bool vec = (HN % VN == 0) && (sp.ld % VN == 0) && (sp.off % VN == 0) && (col0 + HN <= lim);
        for (int s = 0; s < sp.n; s++) {
          T* dst = sp.dst[s] + (long long)row * sp.ld + sp.off + col0;
          if (vec && (reinterpret_cast<uintptr_t>(sp.dst[s]) & 15) == 0) {
#pragma unroll
            for (int j = 0; j < HN; j += VN) *reinterpret_cast<V*>(dst + j) = *reinterpret_cast<V*>(out + j);
          } else {
#pragma unroll
            for (int j = 0; j < HN; j++) if (col0 + j < lim) dst[j] = out[j];
          }
        }
      }
    }
  }
}

}  // namespace guda
