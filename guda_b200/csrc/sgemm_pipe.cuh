// sgemm_pipe.cuh -- the kernel template of sgemm_pipe.cu (see the header comment there), in a header so that the instruction-stream
// explorer (tools/explore_pipe.py: compile-only sweeps read by tools/sass_reuse.py) can instantiate it on its own.
#pragma once
#include "gemm_simt.cuh"
#include "tc_ptx.cuh"

namespace guda {

constexpr int PF_M = 64, PF_K = 16;          // flag granularity: 64 rows of op(A) x 16 k
constexpr int PIPE_ZF_MAX = 2048;            // k-tiles per launch whose flags fit the kernel's shared table

// ---------------------------------------------------------------- main kernel
template <int TM_, int TN_, int LY_, int LX_, int WM_, int WN_, int BK_, int STAGES_, int MINB_, int ORDER_ = 0, bool DRYREG_ = false, int PAIRSEL_ = -1,
          bool ACC64_ = false>
struct PipeCfg {
  static constexpr bool ACC64 = ACC64_;       // accumulators live as packed 64-bit pairs (one asm operand, read-modify-write) instead of float pairs
  static constexpr int ORDER = ORDER_;        // order in which a k step walks its TM x TN outer product (see kstep)
  static constexpr bool DRYREG = DRYREG_;     // timing experiment: fragments stay in registers (no LDS in the k loop); results are garbage
  static constexpr int TM = TM_, TN = TN_, LY = LY_, LX = LX_, WM = WM_, WN = WN_, BK = BK_, STAGES = STAGES_, MINB = MINB_;
  static constexpr int WARPS = WM * WN, THREADS = WARPS * 32;
  static constexpr int BM = WM * TM * LY, BN = WN * TN * LX;
  static constexpr int STAGE_BYTES = BK * (BM + BN) * 4;
  static constexpr int SMEM = STAGES * STAGE_BYTES + 2 * STAGES * 8 + PIPE_ZF_MAX + 128;
  // FFMA2 pairs run along m (PAIR_M) or along n; the other side supplies the broadcast scalar.  Default: pairs along the longer side
  static constexpr bool PAIR_M = PAIRSEL_ < 0 ? (TN <= TM) : (PAIRSEL_ == 1);
  static_assert(LY * LX == 32, "a warp is LY x LX lanes");
  static_assert(TM % 4 == 0 && TN % 4 == 0, "fragments are 128-bit chunks");
  static_assert(BM % PF_M == 0 && BK % PF_K == 0, "flag granularity");
  static_assert(BM <= 256 && BN <= 256, "TMA box");
  static_assert(STAGE_BYTES % 128 == 0, "stage alignment");
};

template <class Cfg, int EPI, bool SCAT>
__global__ void __launch_bounds__(Cfg::THREADS, Cfg::MINB)
sgemm_pipe_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, GemmParams<float> p,
                  const unsigned char* __restrict__ flags, int flag_ld, int lag, int group_rows, ScatterParams<float, SCAT> sp) {
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
    const int tiles_m = p.m / BM, tiles_n = p.n / BN, group = (group_rows / BM) > 0 ? (group_rows / BM) : 1;
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
    pdl_wait();                          // At / flags come from the pre-pass launched just before; everything above overlapped its tail
    const int pre = KT < STAGES ? KT : STAGES;
    for (int s = 0; s < pre; s++) {
      float* sa = reinterpret_cast<float*>(smem_raw + s * Cfg::STAGE_BYTES);
      mbar_arrive_expect_tx(&full[s], Cfg::STAGE_BYTES);
      tma_load_2d(sa, &map_a, &full[s], m0, s * BK);
      tma_load_2d(sa + BK * BM, &map_b, &full[s], n0, s * BK);
    }
  }
  if (tid != 0) pdl_wait();
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
  constexpr int ACC_S = PAIR_M ? TN : TM, ACC_P = PAIR_M ? TM : TN;      // [scalar-side index][paired-side index]
  float acc[Cfg::ACC64 ? 1 : ACC_S][Cfg::ACC64 ? 2 : ACC_P];
  unsigned long long acc2[Cfg::ACC64 ? ACC_S : 1][Cfg::ACC64 ? ACC_P / 2 : 1];
  auto acc_get = [&](int i, int j) -> float {
    const int si = PAIR_M ? j : i, pj = PAIR_M ? i : j;
    if constexpr (Cfg::ACC64) {
      float lo, hi;
      asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc2[si][pj / 2]));
      return (pj & 1) ? hi : lo;
    } else return acc[si][pj];
  };
  auto acc_set = [&](int i, int j, float v) {
    const int si = PAIR_M ? j : i, pj = PAIR_M ? i : j;
    if constexpr (Cfg::ACC64) {
      float lo, hi;
      asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc2[si][pj / 2]));
      if (pj & 1) hi = v; else lo = v;
      asm("mov.b64 %0, {%1, %2};" : "=l"(acc2[si][pj / 2]) : "f"(lo), "f"(hi));
    } else acc[si][pj] = v;
  };
  const int row_base = m0 + wm * (TM * LY) + ly * 4, col_base = n0 + wn * (TN * LX) + lx * 4;
#pragma unroll
  for (int i = 0; i < TM; i++)
#pragma unroll
    for (int j = 0; j < TN; j++) acc_set(i, j, 0.0f);
  if (EPI == EPI_NONE && p.beta != 0.0f) {
#pragma unroll
    for (int i = 0; i < TM; i++) {
      const int row = row_base + (i / 4) * (LY * 4) + (i % 4);
#pragma unroll
      for (int cn = 0; cn < TN / 4; cn++) {
        const float4 v = *reinterpret_cast<const float4*>(p.c + (long long)row * p.ldc + col_base + cn * (LX * 4));
        const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 4; q++) acc_set(i, cn * 4 + q, p.beta == 1.0f ? vv[q] : __fmul_rn(vv[q], p.beta));
      }
    }
  }

  const int a_off = wm * (TM * LY) + ly * 4, b_off = wn * (TN * LX) + lx * 4;
  // one k step of the fast path: fragment loads (128-bit, conflict-free) + TM*TN/2 packed FMAs
  float afx[TM], bfx[TN];
  if constexpr (Cfg::DRYREG) {
#pragma unroll
    for (int i = 0; i < TM; i++) afx[i] = reinterpret_cast<const float*>(smem_raw)[tid + i * 7];
#pragma unroll
    for (int j = 0; j < TN; j++) bfx[j] = reinterpret_cast<const float*>(smem_raw)[tid + 300 + j * 5];
  }
  auto kstep = [&](const float* As, const float* Bs, int kk) {
    float af[TM], bf[TN];
    if constexpr (Cfg::DRYREG) {
#pragma unroll
      for (int i = 0; i < TM; i++) af[i] = afx[i];
#pragma unroll
      for (int j = 0; j < TN; j++) bf[j] = bfx[j];
    } else {
#pragma unroll
      for (int cm = 0; cm < TM / 4; cm++) *reinterpret_cast<float4*>(af + cm * 4) = *reinterpret_cast<const float4*>(As + kk * BM + cm * (LY * 4));
#pragma unroll
      for (int cn = 0; cn < TN / 4; cn++) *reinterpret_cast<float4*>(bf + cn * 4) = *reinterpret_cast<const float4*>(Bs + kk * BN + cn * (LX * 4));
    }
    // Walk order of the outer product.  Every FFMA2 reads an accumulator pair, a fragment pair and a broadcast scalar: five
    // registers, three of them on one side of the even/odd register banks, unless one operand is still in the operand-reuse
    // cache from the previous instruction.  ORDER 0 / 1: plain nested loops; ORDER 2 / 3: boustrophedon -- consecutive FFMA2s
    // always share the pair (or the scalar), so the reuse cache covers the group boundaries too.
    constexpr int NP = (PAIR_M ? TM : TN) / 2, NS = PAIR_M ? TN : TM;      // pairs along the paired side, scalars along the other
    auto one = [&](int pi, int si) {
      if constexpr (Cfg::ACC64) {
        unsigned long long pr;
        if constexpr (PAIR_M) asm("mov.b64 %0, {%1, %2};" : "=l"(pr) : "f"(af[2 * pi]), "f"(af[2 * pi + 1]));
        else asm("mov.b64 %0, {%1, %2};" : "=l"(pr) : "f"(bf[2 * pi]), "f"(bf[2 * pi + 1]));
        asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc2[si][pi]) : "l"(dup_f32(PAIR_M ? bf[si] : af[si])), "l"(pr));
      } else if constexpr (PAIR_M) fma2_f32(acc[si][2 * pi], acc[si][2 * pi + 1], dup_f32(bf[si]), af[2 * pi], af[2 * pi + 1]);
      else fma2_f32(acc[si][2 * pi], acc[si][2 * pi + 1], dup_f32(af[si]), bf[2 * pi], bf[2 * pi + 1]);
    };
    if constexpr (Cfg::ORDER == 0) {
#pragma unroll
      for (int si = 0; si < NS; si++)
#pragma unroll
        for (int pi = 0; pi < NP; pi++) one(pi, si);
    } else if constexpr (Cfg::ORDER == 1) {
#pragma unroll
      for (int pi = 0; pi < NP; pi++)
#pragma unroll
        for (int si = 0; si < NS; si++) one(pi, si);
    } else if constexpr (Cfg::ORDER == 2) {
#pragma unroll
      for (int pi = 0; pi < NP; pi++)
#pragma unroll
        for (int q = 0; q < NS; q++) one(pi, (pi & 1) ? NS - 1 - q : q);
    } else {
#pragma unroll
      for (int si = 0; si < NS; si++)
#pragma unroll
        for (int q = 0; q < NP; q++) one((si & 1) ? NP - 1 - q : q, si);
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
            for (int j = 0; j < TN; j++) acc_set(i, j, __fmaf_rn(af[i], bf[j], acc_get(i, j)));
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

  pdl_trigger();               // the k walk is over: the next kernel on the stream may be scheduled under the epilogue stores
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
        o[q] = apply_epilogue<float, EPI>(acc_get(i, cn * 4 + q), bias);
      }
      const float4 v = make_float4(o[0], o[1], o[2], o[3]);
      if (!SCAT || p.c) *reinterpret_cast<float4*>(p.c + (long long)row * p.ldc + col) = v;
      if constexpr (SCAT) {
        for (int d = 0; d < sp.n; d++) *reinterpret_cast<float4*>(sp.dst[d] + (long long)row * sp.ld + sp.off + col) = v;
      }
    }
  }
}

}  // namespace guda
