// sgemm_pipe.cu -- exact-order SGEMM (NN / TN forms) on a TMA-fed, mbarrier-synchronised shared-memory ring.
//
// Same arithmetic contract as gemm_simt.cuh (compute/sgemm.go:139-163 beta pass, :288-325 the serial FMA forms):
//   acc = beta-scaled c;  for l ascending:  t = fl(alpha * a_il);  if (t != 0) acc = fma(t, b_lj, acc)
// with one thread owning each C element for the whole k walk, so every bit is the reference's.  What changes is how the
// operands reach the FMA pipe:
//   * a pre-pass (sgemm_prepass_kernel) writes t = fl(alpha * op(A)) ONCE, k-major (At[k][m]), and one flag byte per
//     64 x 16 block of t saying whether it holds a zero -- the scaling and the zero test leave the main loop;
//   * the main kernel's operand tiles At[k0:k0+BK][m0:m0+BM] and B[k0:k0+BK][n0:n0+BN] are both "BK rows of contiguous
//     floats": one elected thread issues two cp.async.bulk.tensor (TMA) loads per stage into a STAGES-deep ring, completion
//     on a `full` mbarrier; consumers release a stage on an `empty` mbarrier.  No LDG -> register -> STS staging, no
//     shared-memory transposition (the round-1 kernel's 19 M bank conflicts per launch), no CTA-wide barrier in the k loop;
//   * a warp is a LY x LX grid of lanes (1 x 32: every A fragment load is one warp-wide broadcast wavefront); the
//     micro-tile is TM x TN in 4-wide chunks LY*4 (LX*4) apart so every fragment load is a conflict-free LDS.128;
//     the inner product is packed fma.rn.f32x2 (FFMA2: two IEEE FMAs per issue slot, same bits as two fmaf()).
// A k-tile whose flag is set (a zero in t, or the zero-filled tail of a ragged K) runs the predicated loop instead.
#include "gemm_simt.cuh"
#include "tc_ptx.cuh"
#include <cstdlib>
#include <mutex>

namespace guda {

int tensor_map_2d_f32(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols, uint32_t box_rows);

constexpr int PF_M = 64, PF_K = 16;          // flag granularity: 64 rows of op(A) x 16 k
constexpr int PIPE_ZF_MAX = 2048;            // k-tiles per launch whose flags fit the kernel's shared table

// ---------------------------------------------------------------- pre-pass
// One CTA: a 64 (m) x 64 (k) block of op(A).  TA == false: A is m x k row-major (k contiguous) and the block is transposed
// through shared memory; TA == true: A is stored k x m already.  Requires m % 64 == 0, k % 16 == 0, 16-byte aligned rows.
template <bool TA>
__global__ void __launch_bounds__(256) sgemm_prepass_kernel(const float* __restrict__ a, long long lda, int m, int k, float alpha,
                                                            float* __restrict__ at, long long ldat, unsigned char* __restrict__ flags,
                                                            int flag_ld) {
  __shared__ float tile[64][65];
  __shared__ int zf[4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * 64, k0 = blockIdx.x * 64;
  if (tid < 4) zf[tid] = 0;
  __syncthreads();
  bool z = false;
  if (!TA) {
    const int g = (tid % 16) / 4;                       // this thread's 16-deep k group (the same for all four loads)
#pragma unroll
    for (int it = 0; it < 4; it++) {
      const int v = tid + it * 256, row = v / 16, kc = (v % 16) * 4;
      if (k0 + kc < k) {
        float4 x = *reinterpret_cast<const float4*>(a + (long long)(m0 + row) * lda + k0 + kc);
        x.x = __fmul_rn(alpha, x.x); x.y = __fmul_rn(alpha, x.y); x.z = __fmul_rn(alpha, x.z); x.w = __fmul_rn(alpha, x.w);
        z |= (x.x == 0.0f) | (x.y == 0.0f) | (x.z == 0.0f) | (x.w == 0.0f);
        tile[kc + 0][row] = x.x; tile[kc + 1][row] = x.y; tile[kc + 2][row] = x.z; tile[kc + 3][row] = x.w;
      }
    }
    if (z) zf[g] = 1;
    __syncthreads();
#pragma unroll
    for (int it = 0; it < 4; it++) {
      const int v = tid + it * 256, kk = v / 16, mc = (v % 16) * 4;
      if (k0 + kk < k) {
        float4 o = make_float4(tile[kk][mc], tile[kk][mc + 1], tile[kk][mc + 2], tile[kk][mc + 3]);
        *reinterpret_cast<float4*>(at + (long long)(k0 + kk) * ldat + m0 + mc) = o;
      }
    }
  } else {
#pragma unroll
    for (int it = 0; it < 4; it++) {
      const int v = tid + it * 256, kk = v / 16, mc = (v % 16) * 4;
      if (k0 + kk < k) {
        float4 x = *reinterpret_cast<const float4*>(a + (long long)(k0 + kk) * lda + m0 + mc);
        x.x = __fmul_rn(alpha, x.x); x.y = __fmul_rn(alpha, x.y); x.z = __fmul_rn(alpha, x.z); x.w = __fmul_rn(alpha, x.w);
        if ((x.x == 0.0f) | (x.y == 0.0f) | (x.z == 0.0f) | (x.w == 0.0f)) zf[kk / 16] = 1;
        *reinterpret_cast<float4*>(at + (long long)(k0 + kk) * ldat + m0 + mc) = x;
      }
    }
    __syncthreads();
  }
  if (tid < 4 && k0 + tid * 16 < k) flags[(long long)blockIdx.y * flag_ld + blockIdx.x * 4 + tid] = (unsigned char)zf[tid];
}

// ---------------------------------------------------------------- main kernel
template <int TM_, int TN_, int LY_, int LX_, int WM_, int WN_, int BK_, int STAGES_, int MINB_>
struct PipeCfg {
  static constexpr int TM = TM_, TN = TN_, LY = LY_, LX = LX_, WM = WM_, WN = WN_, BK = BK_, STAGES = STAGES_, MINB = MINB_;
  static constexpr int WARPS = WM * WN, THREADS = WARPS * 32;
  static constexpr int BM = WM * TM * LY, BN = WN * TN * LX;
  static constexpr int STAGE_BYTES = BK * (BM + BN) * 4;
  static constexpr int SMEM = STAGES * STAGE_BYTES + 2 * STAGES * 8 + PIPE_ZF_MAX + 128;
  static constexpr bool PAIR_M = TN <= TM;          // duplicate the operand of the shorter micro-tile side
  static_assert(LY * LX == 32, "a warp is LY x LX lanes");
  static_assert(TM % 4 == 0 && TN % 4 == 0, "fragments are 128-bit chunks");
  static_assert(BM % PF_M == 0 && BK % PF_K == 0, "flag granularity");
  static_assert(BM <= 256 && BN <= 256, "TMA box");
  static_assert(STAGE_BYTES % 128 == 0, "stage alignment");
};

template <class Cfg, int EPI, bool SCAT>
__global__ void __launch_bounds__(Cfg::THREADS, Cfg::MINB)
sgemm_pipe_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, GemmParams<float> p,
                  const unsigned char* __restrict__ flags, int flag_ld, int lag, ScatterParams<float, SCAT> sp) {
  using namespace tc;
  constexpr int TM = Cfg::TM, TN = Cfg::TN, LY = Cfg::LY, LX = Cfg::LX, BM = Cfg::BM, BN = Cfg::BN, BK = Cfg::BK, STAGES = Cfg::STAGES;
  constexpr bool PAIR_M = Cfg::PAIR_M;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  unsigned char* zf = reinterpret_cast<unsigned char*>(empty + STAGES);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int wm = warp / Cfg::WN, wn = warp % Cfg::WN, ly = lane / LX, lx = lane % LX;
  // CTA -> tile: groups of row blocks sweep N so that co-resident CTAs share A row panels and B column panels in L2
  int tile_m, tile_n;
  {
    const int tiles_m = p.m / BM, tiles_n = p.n / BN, group = (2048 / BM) > 0 ? (2048 / BM) : 1;
    const int t = blockIdx.x;
    const int per_group = group * tiles_n;
    const int gi = t / per_group, r = t - gi * per_group;
    const int rows = min(group, tiles_m - gi * group);
    tile_m = gi * group + r % rows;
    tile_n = r / rows;
  }
  const int m0 = tile_m * BM, n0 = tile_n * BN;
  const int KT = p.k / BK;

  // thread 0 arms the ring and requests the first STAGES k-tiles before anything else, so the first TMA round trip runs
  // under the flag-table load and the accumulator initialisation of the other threads
  if (tid == 0) {
    for (int s = 0; s < STAGES; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], Cfg::WARPS); }
    fence_barrier_init();
    prefetch_tmap(&map_a); prefetch_tmap(&map_b);
    const int pre = KT < STAGES ? KT : STAGES;
    for (int s = 0; s < pre; s++) {
      float* sa = reinterpret_cast<float*>(smem_raw + s * Cfg::STAGE_BYTES);
      mbar_arrive_expect_tx(&full[s], Cfg::STAGE_BYTES);
      tma_load_2d(sa, &map_a, &full[s], m0, s * BK);
      tma_load_2d(sa + BK * BM, &map_b, &full[s], n0, s * BK);
    }
  }
  for (int kt = tid; kt < KT; kt += Cfg::THREADS) {
    unsigned char f = 0;
#pragma unroll
    for (int rb = 0; rb < BM / PF_M; rb++)
#pragma unroll
      for (int q = 0; q < BK / PF_K; q++) f |= flags[(long long)(m0 / PF_M + rb) * flag_ld + kt * (BK / PF_K) + q];
    zf[kt] = f;
  }
  __syncthreads();

  // accumulators: PAIR_M keeps (i, i+1) adjacent so a pair is one 64-bit register
  float acc[PAIR_M ? TN : TM][PAIR_M ? TM : TN];
#define ACC(i, j) (PAIR_M ? acc[j][i] : acc[i][j])
  const int row_base = m0 + wm * (TM * LY) + ly * 4, col_base = n0 + wn * (TN * LX) + lx * 4;
#pragma unroll
  for (int i = 0; i < TM; i++)
#pragma unroll
    for (int j = 0; j < TN; j++) ACC(i, j) = 0.0f;
  if (EPI == EPI_NONE && p.beta != 0.0f) {
#pragma unroll
    for (int i = 0; i < TM; i++) {
      const int row = row_base + (i / 4) * (LY * 4) + (i % 4);
#pragma unroll
      for (int cn = 0; cn < TN / 4; cn++) {
        const float4 v = *reinterpret_cast<const float4*>(p.c + (long long)row * p.ldc + col_base + cn * (LX * 4));
        const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 4; q++) ACC(i, cn * 4 + q) = p.beta == 1.0f ? vv[q] : __fmul_rn(vv[q], p.beta);
      }
    }
  }

  const int a_off = wm * (TM * LY) + ly * 4, b_off = wn * (TN * LX) + lx * 4;
  // one k step of the fast path: fragment loads (128-bit, conflict-free) + TM*TN/2 packed FMAs
  auto kstep = [&](const float* As, const float* Bs, int kk) {
    float af[TM], bf[TN];
#pragma unroll
    for (int cm = 0; cm < TM / 4; cm++) *reinterpret_cast<float4*>(af + cm * 4) = *reinterpret_cast<const float4*>(As + kk * BM + cm * (LY * 4));
#pragma unroll
    for (int cn = 0; cn < TN / 4; cn++) *reinterpret_cast<float4*>(bf + cn * 4) = *reinterpret_cast<const float4*>(Bs + kk * BN + cn * (LX * 4));
    if constexpr (PAIR_M) {
#pragma unroll
      for (int j = 0; j < TN; j++) {
        const unsigned long long bb = dup_f32(bf[j]);
#pragma unroll
        for (int i = 0; i < TM; i += 2) fma2_f32(acc[j][i], acc[j][i + 1], bb, af[i], af[i + 1]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < TM; i++) {
        const unsigned long long aa = dup_f32(af[i]);
#pragma unroll
        for (int j = 0; j < TN; j += 2) fma2_f32(acc[i][j], acc[i][j + 1], aa, bf[j], bf[j + 1]);
      }
    }
  };
  unsigned char zcur = KT > 0 ? zf[0] : 0;
  // s / phase: ring slot and mbarrier parity of the current k-tile, carried in registers (no division in the loop).
  // `ready`: the full barrier of the NEXT k-tile is tested (non-blocking mbarrier.test_wait) in the middle of the current
  // stage's arithmetic, so at the stage boundary a warp normally falls straight through instead of paying the
  // try_wait round trip (ncu: 4.6 % of all warp samples sat on that branch in the first version of this kernel).
  int s = 0;
  uint32_t phase = 0, ready = 0;
  const bool dry = lag >= 100;      // timing experiment only (guda_set_sgemm_pipe lag >= 100): no waits, no refills, results are garbage
  if (dry) lag -= 100;
  for (int kt = 0; kt < KT; kt++) {
    const float* As = reinterpret_cast<const float*>(smem_raw + s * Cfg::STAGE_BYTES) + a_off;
    const float* Bs = reinterpret_cast<const float*>(smem_raw + s * Cfg::STAGE_BYTES) + BK * BM + b_off;
    if (!ready && !dry) mbar_wait(&full[s], phase);
    const int s1 = (s + 1 == STAGES) ? 0 : s + 1;
    const uint32_t phase1 = (s + 1 == STAGES) ? phase ^ 1u : phase;
    unsigned char znext = 0;
    if (!zcur) {
#pragma unroll
      for (int kk = 0; kk < BK / 2; kk++) kstep(As, Bs, kk);
      if (kt + 1 < KT) { ready = dry ? 1u : mbar_test_wait_nb(&full[s1], phase1); znext = zf[kt + 1]; }
#pragma unroll
      for (int kk = BK / 2; kk < BK; kk++) kstep(As, Bs, kk);
    } else {
#pragma unroll 2
      for (int kk = 0; kk < BK; kk++) {
        float af[TM], bf[TN];
#pragma unroll
        for (int cm = 0; cm < TM / 4; cm++) *reinterpret_cast<float4*>(af + cm * 4) = *reinterpret_cast<const float4*>(As + kk * BM + cm * (LY * 4));
#pragma unroll
        for (int cn = 0; cn < TN / 4; cn++) *reinterpret_cast<float4*>(bf + cn * 4) = *reinterpret_cast<const float4*>(Bs + kk * BN + cn * (LX * 4));
#pragma unroll
        for (int i = 0; i < TM; i++) {
          if (af[i] != 0.0f) {                                  // compute/sgemm.go:301 `if tmp != 0` (true for NaN)
#pragma unroll
            for (int j = 0; j < TN; j++) ACC(i, j) = __fmaf_rn(af[i], bf[j], ACC(i, j));
          }
        }
      }
      ready = 0;
      if (kt + 1 < KT) znext = zf[kt + 1];
    }
    zcur = znext;
    __syncwarp();
    if (lane == 0 && !dry) {
      mbar_arrive(&empty[s]);
      // the refill of the slot every warp released `lag` iterations ago (k-tile kr + STAGES) rotates over the warps, so no
      // single warp carries the producer's instructions on its critical path
      const int kr = kt - lag;
      if (kr >= 0 && kr + STAGES < KT && (kt % Cfg::WARPS) == warp) {
        int sr = s - lag; if (sr < 0) sr += STAGES;
        const uint32_t pr = (s - lag < 0) ? phase ^ 1u : phase;
        mbar_wait(&empty[sr], pr);
        float* sa = reinterpret_cast<float*>(smem_raw + sr * Cfg::STAGE_BYTES);
        mbar_arrive_expect_tx(&full[sr], Cfg::STAGE_BYTES);
        tma_load_2d(sa, &map_a, &full[sr], m0, (kr + STAGES) * BK);
        tma_load_2d(sa + BK * BM, &map_b, &full[sr], n0, (kr + STAGES) * BK);
      }
    }
    s = s1; phase = phase1;
  }
  // (every refill is issued inside the loop: k-tile kr + STAGES < KT is requested at iteration kr + lag <= KT - 1 for lag < STAGES)

  // ---- epilogue in registers, single 128-bit store per chunk
#pragma unroll
  for (int i = 0; i < TM; i++) {
    const int row = row_base + (i / 4) * (LY * 4) + (i % 4);
#pragma unroll
    for (int cn = 0; cn < TN / 4; cn++) {
      const int col = col_base + cn * (LX * 4);
      float o[4];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        float bias = 0.0f;
        if (EPI != EPI_NONE) bias = p.bias[col + q];
        o[q] = apply_epilogue<float, EPI>(ACC(i, cn * 4 + q), bias);
      }
      const float4 v = make_float4(o[0], o[1], o[2], o[3]);
      if (!SCAT || p.c) *reinterpret_cast<float4*>(p.c + (long long)row * p.ldc + col) = v;
      if constexpr (SCAT) {
        for (int d = 0; d < sp.n; d++) *reinterpret_cast<float4*>(sp.dst[d] + (long long)row * sp.ld + sp.off + col) = v;
      }
    }
  }
#undef ACC
}

// ---------------------------------------------------------------- host side
//                    TM  TN  LY  LX  WM WN  BK ST MINB
using Pipe1  = PipeCfg< 8,  8,  2, 16,  8, 1, 16, 4, 2>;    // 128 x 128, 256 threads, 2 CTAs/SM (the round-1 shape on the new pipeline)
using Pipe2  = PipeCfg< 8,  8,  1, 32,  8, 1, 16, 4, 2>;    //  64 x 256, warp = one lane row: A fragments are warp-wide broadcasts
using Pipe3  = PipeCfg<16,  8,  1, 32,  8, 1, 16, 4, 1>;    // 128 x 256, 16 x 8 micro-tile, 1 CTA/SM
using Pipe4  = PipeCfg<16,  8,  2, 16,  4, 1, 16, 4, 2>;    // 128 x 128, 128 threads, 16 x 8 micro-tile, 2 CTAs/SM
using Pipe5  = PipeCfg<16,  8,  2, 16,  4, 1, 32, 3, 2>;    // Pipe4 with 32-deep stages
using Pipe6  = PipeCfg< 8,  4,  1, 32,  8, 1, 16, 4, 2>;    //  64 x 128 (mid sizes), 8 x 4 micro-tile
using Pipe7  = PipeCfg< 8,  8,  2, 16,  4, 1, 16, 4, 2>;    //  64 x 128, 128 threads, 8 x 8 micro-tile
using Pipe8  = PipeCfg< 8,  8,  2, 16,  8, 1, 32, 3, 2>;    // Pipe1 with 32-deep stages
using Pipe9  = PipeCfg<16,  8,  1, 32,  4, 1, 16, 4, 2>;    //  64 x 256, 128 threads, 16 x 8 micro-tile on one lane row
using Pipe10 = PipeCfg< 8,  8,  2, 16,  4, 1, 32, 3, 2>;    // Pipe7 with 32-deep stages
using Pipe11 = PipeCfg<16,  8,  1, 32,  4, 1, 16, 5, 2>;    // Pipe9 with a 5-deep ring
using Pipe12 = PipeCfg<16,  8,  2, 16,  4, 1, 16, 6, 2>;    // Pipe4 with a 6-deep ring
using Pipe13 = PipeCfg<16,  8,  1, 32,  4, 1, 32, 2, 2>;    // Pipe9 with 32-deep stages, 2-deep ring
using Pipe14 = PipeCfg< 8,  8,  2, 16,  4, 1, 64, 2, 2>;    // Pipe7 with 64-deep stages (1024^3: 16 stage boundaries)
constexpr int PIPE_NCFG = 14;

static int g_pipe_cfg = [] { const char* e = getenv("GUDA_SGEMM_PIPE"); return e ? atoi(e) : -1; }();   // -1 auto, 0 off, 1.. forced config
static int g_pipe_lag = [] { const char* e = getenv("GUDA_SGEMM_PIPE_LAG"); return e ? atoi(e) : 1; }();

template <class Cfg, int EPI, bool SCAT>
static int launch_pipe_cfg(const GemmParams<float>& p, const float* at, long long ldat, const unsigned char* flags, int flag_ld,
                           const ScatterParams<float, SCAT>& sp, cudaStream_t s) {
  CUtensorMap ma, mb;
  int rc;
  if ((rc = tensor_map_2d_f32(&ma, at, (uint64_t)p.k, (uint64_t)p.m, (uint64_t)ldat, Cfg::BM, Cfg::BK))) return rc;
  if ((rc = tensor_map_2d_f32(&mb, p.b, (uint64_t)p.k, (uint64_t)p.n, (uint64_t)p.ldb, Cfg::BN, Cfg::BK))) return rc;
  auto kern = sgemm_pipe_kernel<Cfg, EPI, SCAT>;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [&] { attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM); });
  if (attr_err != cudaSuccess) return set_error(GUDA_ERR_EXECUTION, "sgemm_pipe: cannot reserve %d bytes of shared memory", Cfg::SMEM);
  const long long tiles = (long long)(p.m / Cfg::BM) * (p.n / Cfg::BN);
  int lag = g_pipe_lag < 0 ? 0 : (g_pipe_lag >= 100 ? 101 : (g_pipe_lag > Cfg::STAGES - 1 ? Cfg::STAGES - 1 : g_pipe_lag));
  kern<<<(unsigned)tiles, Cfg::THREADS, Cfg::SMEM, s>>>(ma, mb, p, flags, flag_ld, lag, sp);
  return check_launch("sgemm_pipe");
}

template <class Cfg>
static bool pipe_fits(int m, int n, int k) { return m % Cfg::BM == 0 && n % Cfg::BN == 0 && k % Cfg::BK == 0; }

template <int EPI, bool SCAT>
static int launch_pipe(int cfg, const GemmParams<float>& p, const float* at, long long ldat, const unsigned char* flags, int flag_ld,
                       const ScatterParams<float, SCAT>& sp, cudaStream_t s) {
  switch (cfg) {
#define GUDA_PIPE_CASE(ID, CFG) case ID: return launch_pipe_cfg<CFG, EPI, SCAT>(p, at, ldat, flags, flag_ld, sp, s);
    GUDA_PIPE_CASE(1, Pipe1) GUDA_PIPE_CASE(2, Pipe2) GUDA_PIPE_CASE(3, Pipe3) GUDA_PIPE_CASE(4, Pipe4) GUDA_PIPE_CASE(5, Pipe5)
    GUDA_PIPE_CASE(6, Pipe6) GUDA_PIPE_CASE(7, Pipe7) GUDA_PIPE_CASE(8, Pipe8) GUDA_PIPE_CASE(9, Pipe9) GUDA_PIPE_CASE(10, Pipe10)
    GUDA_PIPE_CASE(11, Pipe11) GUDA_PIPE_CASE(12, Pipe12) GUDA_PIPE_CASE(13, Pipe13) GUDA_PIPE_CASE(14, Pipe14)
#undef GUDA_PIPE_CASE
  }
  return set_error(GUDA_ERR_INVALID_ARG, "sgemm_pipe: unknown configuration %d", cfg);
}

static bool pipe_cfg_fits(int cfg, int m, int n, int k) {
  switch (cfg) {
#define GUDA_PIPE_FIT(ID) case ID: return pipe_fits<Pipe##ID>(m, n, k);
    GUDA_PIPE_FIT(1) GUDA_PIPE_FIT(2) GUDA_PIPE_FIT(3) GUDA_PIPE_FIT(4) GUDA_PIPE_FIT(5) GUDA_PIPE_FIT(6) GUDA_PIPE_FIT(7)
    GUDA_PIPE_FIT(8) GUDA_PIPE_FIT(9) GUDA_PIPE_FIT(10) GUDA_PIPE_FIT(11) GUDA_PIPE_FIT(12) GUDA_PIPE_FIT(13) GUDA_PIPE_FIT(14)
#undef GUDA_PIPE_FIT
  }
  return false;
}

// Returns GUDA_OK / an error if the pipelined route took the call, -1 if the call is not eligible (the caller then uses
// the register-staged kernel of gemm_simt.cuh).  Forms: NN and TN (B must be k x n row-major); fp32 FMA chain only.
int sgemm_pipe_try(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb, float beta, float* c,
                   int ldc, const float* bias, int epi, const ScatterArgs* scatter, cudaStream_t s) {
  int cfg = g_pipe_cfg;
  if (cfg == 0 || tb || k <= 0) return -1;
  if (cfg < 0) return -1;                        // auto selection is wired in once the configurations are measured
  if (!pipe_cfg_fits(cfg, m, n, k)) return -1;
  auto al16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  if (!al16(a) || !al16(b) || (c && !al16(c)) || lda % 4 || ldb % 4 || ldc % 4) return -1;
  if (scatter) {
    if (scatter->ld % 4 || scatter->off % 4) return -1;
    for (int i = 0; i < scatter->n; i++) if (!al16(scatter->dst[i])) return -1;
  }
  constexpr int KC = 8192;
  const int kc_max = (epi == EPI_NONE && k >= 2 * KC) ? KC : k;       // the fused epilogues run on one unchunked chain
  if (kc_max / 16 > PIPE_ZF_MAX) return -1;
  PoolTemp at, fl;
  int rc;
  const int flag_ld = (kc_max + PF_K - 1) / PF_K;
  if ((rc = at.alloc((size_t)kc_max * m * sizeof(float), s))) return rc;
  if ((rc = fl.alloc((size_t)(m / PF_M) * flag_ld, s))) return rc;
  for (int k0 = 0; k0 < k; k0 += kc_max) {
    const int kc = k - k0 < kc_max ? k - k0 : kc_max;
    const bool last = k0 + kc >= k;
    dim3 pg((kc + 63) / 64, m / 64);
    if (ta) sgemm_prepass_kernel<true><<<pg, 256, 0, s>>>(a + (long long)k0 * lda, lda, m, kc, alpha, (float*)at.p, m, (unsigned char*)fl.p, flag_ld);
    else    sgemm_prepass_kernel<false><<<pg, 256, 0, s>>>(a + k0, lda, m, kc, alpha, (float*)at.p, m, (unsigned char*)fl.p, flag_ld);
    if ((rc = check_launch("sgemm_prepass"))) return rc;
    GemmParams<float> p{m, n, kc, alpha, k0 == 0 ? beta : 1.0f, nullptr, 0, b + (long long)k0 * ldb, ldb, c, ldc, bias, m, n};
    if (scatter && last) {
      ScatterParams<float, true> sp;
      sp.n = scatter->n;
      for (int i = 0; i < GUDA_MAX_SCATTER; i++) sp.dst[i] = i < scatter->n ? scatter->dst[i] : nullptr;
      sp.ld = scatter->ld; sp.off = scatter->off;
      rc = launch_pipe<EPI_NONE, true>(cfg, p, (const float*)at.p, m, (const unsigned char*)fl.p, flag_ld, sp, s);
    } else if (!last || epi == EPI_NONE) {
      if (!last && (scatter && !c)) return set_error(GUDA_ERR_INVALID_ARG, "sgemm_scatter: a K-chunked call needs a local C block");
      rc = launch_pipe<EPI_NONE, false>(cfg, p, (const float*)at.p, m, (const unsigned char*)fl.p, flag_ld, {}, s);
    } else if (epi == EPI_BIAS_RELU) {
      rc = launch_pipe<EPI_BIAS_RELU, false>(cfg, p, (const float*)at.p, m, (const unsigned char*)fl.p, flag_ld, {}, s);
    } else {
      rc = launch_pipe<EPI_BIAS_GELU, false>(cfg, p, (const float*)at.p, m, (const unsigned char*)fl.p, flag_ld, {}, s);
    }
    if (rc) return rc;
  }
  return GUDA_OK;
}

}  // namespace guda

// A/B and test knob: cfg -1 = automatic choice, 0 = pipeline off (register-staged kernel), 1.. = force a configuration
// (the call falls back to the register-staged kernel when the shape does not fit it); lag = refill distance of the producer.
extern "C" int guda_set_sgemm_pipe(int cfg, int lag) {
  if (cfg < -1 || cfg > guda::PIPE_NCFG) return guda::set_error(GUDA_ERR_INVALID_ARG, "unknown pipeline configuration %d", cfg);
  guda::g_pipe_cfg = cfg;
  if (lag >= 0) guda::g_pipe_lag = lag;
  return GUDA_OK;
}
