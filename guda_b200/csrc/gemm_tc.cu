// gemm_tc.cu -- Blackwell tensor-core GEMM: TMA-fed tcgen05.mma with fp32 accumulators in TMEM.
//
// One persistent, warp-specialised kernel (1 CTA per SM, 192 threads):
//   warp 0      TMA producer: cp.async.bulk.tensor tiles of A (K-major, rows of 128 B) and B (row-major [K][N] as
//               stored by GUDA = "MN-major": boxes of BK x 128 B) into a ring of SWIZZLE_128B shared-memory stages
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma (M=128, N=BN) per 32-byte K slice; tcgen05.commit
//               releases smem stages and signals the epilogue; owns the TMEM allocation (2 accumulator stages)
//   warps 2..5  epilogue: tcgen05.ld 32 lanes x 32 columns -> registers -> alpha/beta, bias + ReLU/GELU in registers
//               -> ONE store of C (subsystem 4: no second pass over C)
// Instantiations:
//   16-bit (bf16 / fp16 storage, fp32 accumulate)    replaces bfloat16.go:104-167, float16.go:201-244
//   3xTF32 SGEMM: A = Ah + Al, B = Bh + Bl (fp32 planes produced by a split pre-pass); per K slice
//                 D += Al*Bh; D += Ah*Bl; D += Ah*Bh   (kind::tf32, fp32 TMEM accumulation)      GUDA_SGEMM_3XTF32
// Shared-memory operand layout and the 64-bit descriptors follow the canonical UMMA layouts
// (K-major: ((8,n),2):((8,SBO),1); MN-major: ((8,n),(8,k)):((1,LBO),(8,SBO)), units of 16 bytes).
#include "common.cuh"
#include "gemm_simt.cuh"   // epilogue maths (apply_epilogue / gelu_f32_exact)
#include "tc_ptx.cuh"
#include <cudaTypedefs.h>
#include <cuda_fp16.h>
#include <mutex>

namespace guda {

using namespace tc;

enum { OUT_BF16 = 0, OUT_F16 = 1, OUT_F32 = 2 };

struct TcParams {
  int m, n, k;
  float alpha, beta;
  void* c; long long ldc;
  const float* bias;
  int epi;
  int tiles_m, tiles_n;
};

template <int EB, int NSPLIT, int BN>
struct TcCfg {
  static constexpr int BM = 128;
  static constexpr int BK = 128 / EB;                 // one 128-byte swizzle row of K per stage row
  static constexpr int UMMA_K = 32 / EB;              // 32 bytes of K per tcgen05.mma
  static constexpr int A_TILE = BM * 128;             // bytes
  static constexpr int B_BOX_N = 128 / EB;            // elements of N per TMA box (128 bytes)
  static constexpr int B_BOX = BK * 128;              // bytes per box
  static constexpr int B_TILE = (BN / B_BOX_N) * B_BOX;
  static constexpr int STAGE = NSPLIT * (A_TILE + B_TILE);
  static constexpr int STAGES = (200 * 1024) / STAGE;
  static constexpr int ACC_STAGES = 2;
  static constexpr int TMEM_COLS = ACC_STAGES * BN;   // 512 (BN=256) or 256 (BN=128): powers of two
  static constexpr int SMEM = STAGES * STAGE + 1024 /*align*/ + 256 /*barriers*/;
  static_assert(STAGES >= 2, "pipeline too shallow");
};

__device__ __forceinline__ uint16_t tc_to_bf16(float f) {     // bfloat16.go:12-23
  uint32_t bits = __float_as_uint(f);
  if ((bits & 0x7FFFu) == 0x7FFFu && ((bits >> 16) & 1u)) bits += 0x10000u;
  return (uint16_t)(bits >> 16);
}
__device__ __forceinline__ float tc_from_bf16(uint16_t b) { return __uint_as_float((uint32_t)b << 16); }
__device__ __forceinline__ uint16_t tc_to_f16(float f) {      // float16.go:53-87 (truncate, flush, saturate to Inf)
  uint32_t bits = __float_as_uint(f);
  uint32_t sign = (bits >> 16) & 0x8000u, exponent = (bits >> 23) & 0xFFu, mant = bits & 0x7FFFFFu;
  if (exponent == 0xFFu) return (uint16_t)(mant == 0 ? (sign | 0x7C00u) : (sign | 0x7C00u | (mant >> 13)));
  int e = (int)exponent - 127 + 15;
  if (e <= 0) return (uint16_t)sign;
  if (e >= 0x1F) return (uint16_t)(sign | 0x7C00u);
  return (uint16_t)(sign | ((uint32_t)e << 10) | (mant >> 13));
}
__device__ __forceinline__ float tc_from_f16(uint16_t h) { return __half2float(__ushort_as_half(h)); }

// GELU_tanh for the 16-bit tensor route.  With e = exp(2u), 0.5*x*(1 + tanh(u)) = x - x/(e + 1): one ex2.approx and one
// rcp.approx per element, no saturation branch (e = inf gives x, e = 0 gives 0).  Absolute error ~2e-7, three orders
// below the bf16/fp16 output spacing; the fp32 routes keep the reference's exact geluFloat32.
__device__ __forceinline__ float gelu_fast(float x) {
  const float x2 = __fmul_rn(x, x);
  const float inner = __fmaf_rn(__fmul_rn(0.044715f, x2), x, x);
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(__fmul_rn(2.3022082f, inner)));     // 2 * sqrt(2/pi) * log2(e)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(__fadd_rn(e, 1.0f)));
  return __fmaf_rn(-x, r, x);
}

// bias[col0 .. col0+32) -> registers.  All lanes of a warp read the same addresses (one broadcast transaction per load).
__device__ __forceinline__ void load_bias32(const float* __restrict__ bias, int col0, int n, float (&b)[32]) {
  if (col0 + 32 <= n && ((reinterpret_cast<uintptr_t>(bias + col0) & 15) == 0)) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(bias + col0 + j));
      b[j] = t.x; b[j + 1] = t.y; b[j + 2] = t.z; b[j + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; j++) b[j] = (col0 + j < n) ? __ldg(bias + col0 + j) : 0.0f;
  }
}

// One thread's 32 consecutive columns of one C row: beta*C, bias + activation in registers, one store.
template <int OUT, int EPI>
__device__ __forceinline__ void store_chunk(const TcParams& p, float (&v)[32], int row, int col0) {
  const bool full = (col0 + 32 <= p.n);
  float bias[32];
  if (EPI != EPI_NONE) load_bias32(p.bias, col0, p.n, bias);
  if (OUT == OUT_F32) {
    float* dst = reinterpret_cast<float*>(p.c) + (long long)row * p.ldc + col0;
    if (p.beta != 0.0f) {
#pragma unroll
      for (int j = 0; j < 32; j++) if (full || col0 + j < p.n) v[j] = __fadd_rn(v[j], __fmul_rn(p.beta, dst[j]));
    }
    if (EPI != EPI_NONE) {
#pragma unroll
      for (int j = 0; j < 32; j++) v[j] = apply_epilogue<float, EPI>(v[j], bias[j]);
    }
    if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; j++) if (col0 + j < p.n) dst[j] = v[j];
    }
  } else {
    uint16_t* dst = reinterpret_cast<uint16_t*>(p.c) + (long long)row * p.ldc + col0;
    if (p.beta != 0.0f) {
#pragma unroll
      for (int j = 0; j < 32; j++)
        if (full || col0 + j < p.n) v[j] = __fadd_rn(v[j], __fmul_rn(p.beta, OUT == OUT_BF16 ? tc_from_bf16(dst[j]) : tc_from_f16(dst[j])));
    }
    if (EPI == EPI_BIAS_RELU) {
#pragma unroll
      for (int j = 0; j < 32; j++) v[j] = apply_epilogue<float, EPI_BIAS_RELU>(v[j], bias[j]);
    } else if (EPI == EPI_BIAS_GELU) {
#pragma unroll
      for (int j = 0; j < 32; j++) v[j] = gelu_fast(__fadd_rn(v[j], bias[j]));
    }
    uint32_t packed[16];
#pragma unroll
    for (int j = 0; j < 16; j++) {
      uint32_t lo = OUT == OUT_BF16 ? tc_to_bf16(v[2 * j]) : tc_to_f16(v[2 * j]);
      uint32_t hi = OUT == OUT_BF16 ? tc_to_bf16(v[2 * j + 1]) : tc_to_f16(v[2 * j + 1]);
      packed[j] = lo | (hi << 16);
    }
    if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
      for (int j = 0; j < 16; j += 4) *reinterpret_cast<uint4*>(dst + 2 * j) = make_uint4(packed[j], packed[j + 1], packed[j + 2], packed[j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; j++) if (col0 + j < p.n) dst[j] = (uint16_t)(packed[j >> 1] >> ((j & 1) * 16));
    }
  }
}

template <int EB, int NSPLIT, int BN, int OUT, int EPI>
__global__ void __launch_bounds__(192, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a0, const __grid_constant__ CUtensorMap map_a1,
               const __grid_constant__ CUtensorMap map_b0, const __grid_constant__ CUtensorMap map_b1, TcParams p) {
  using Cfg = TcCfg<EB, NSPLIT, BN>;
  constexpr int BM = Cfg::BM, BK = Cfg::BK, STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + Cfg::ACC_STAGES;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + Cfg::ACC_STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int kblocks = (p.k + BK - 1) / BK;
  // The tensor core adds into TMEM with truncation, so a long-K fp32 sum drifts (measured: ~K/8 ulps for 3xTF32).
  // The 3xTF32 route therefore accumulates at most CHUNK_K = 256 of K in TMEM and promotes each chunk into fp32
  // registers (round-to-nearest adds) in the epilogue warps, as block-scaled FP8 GEMMs do; 16-bit outputs do not need it.
  const int chunk_blocks = (NSPLIT == 2) ? 8 : (kblocks > 0 ? kblocks : 1);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_a0); prefetch_tmap(&map_b0);
    if (NSPLIT == 2) { prefetch_tmap(&map_a1); prefetch_tmap(&map_b1); }
    for (int s = 0; s < STAGES; s++) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < Cfg::ACC_STAGES; s++) { mbar_init(&tmem_full[s], 1); mbar_init(&tmem_empty[s], 4); }
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(tmem_ptr, Cfg::TMEM_COLS); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  // tile order: groups of 8 M-blocks sweep N, so concurrently resident CTAs share A rows and B columns in L2
  auto tile_coord = [&](int t, int& tm, int& tn) {
    const int group = 8;
    int per_group = group * p.tiles_n;
    int g = t / per_group, r = t % per_group;
    int rows = min(group, p.tiles_m - g * group);
    tm = g * group + r % rows;
    tn = r / rows;
  };

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        int tm, tn; tile_coord(t, tm, tn);
        for (int kb = 0; kb < kblocks; kb++) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* st = smem + stage * Cfg::STAGE;
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE);
          tma_load_2d(st, &map_a0, &full_bar[stage], kb * BK, tm * BM);
          if (NSPLIT == 2) tma_load_2d(st + Cfg::A_TILE, &map_a1, &full_bar[stage], kb * BK, tm * BM);
          uint8_t* sb = st + NSPLIT * Cfg::A_TILE;
#pragma unroll
          for (int j = 0; j < BN / Cfg::B_BOX_N; j++) {
            tma_load_2d(sb + j * Cfg::B_BOX, &map_b0, &full_bar[stage], tn * BN + j * Cfg::B_BOX_N, kb * BK);
            if (NSPLIT == 2) tma_load_2d(sb + Cfg::B_TILE + j * Cfg::B_BOX, &map_b1, &full_bar[stage], tn * BN + j * Cfg::B_BOX_N, kb * BK);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(EB == 2 ? (OUT == OUT_F16 ? 0u : 1u) : 2u, /*a K-major*/ 0u, /*b MN-major*/ 1u, BM, BN);
      int stage = 0; uint32_t phase = 0;
      int ci = 0;                                             // running accumulation-chunk index (2 TMEM stages)
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        for (int kb0 = 0; kb0 < kblocks; kb0 += chunk_blocks, ci++) {
          const int as = ci & 1; const uint32_t aphase = (ci >> 1) & 1;
          mbar_wait(&tmem_empty[as], aphase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * BN;
          const int kb1 = min(kb0 + chunk_blocks, kblocks);
          for (int kb = kb0; kb < kb1; kb++) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint32_t a0 = smem_u32(smem + stage * Cfg::STAGE);
            const uint32_t b0 = a0 + NSPLIT * Cfg::A_TILE;
#pragma unroll
            for (int ks = 0; ks < BK / Cfg::UMMA_K; ks++) {
              // A: K-major, 128-byte rows, 8-row groups 1024 B apart; K slice = +32 B inside the swizzled row
              // B: MN-major, boxes (LBO) of BK rows x 128 B, row groups (SBO) apart; K slice = UMMA_K rows
              // (fp32/tf32 MN-major operands only exist as SWIZZLE_128B_BASE32B: 32-byte swizzle atoms, 4-row groups 512 B apart)
              constexpr uint32_t B_LAYOUT = EB == 4 ? 1u : 2u, B_SBO = EB == 4 ? 512u : 1024u;
              const uint64_t da = make_smem_desc(a0 + ks * 32, 16, 1024);
              const uint64_t db = make_smem_desc(b0 + ks * Cfg::UMMA_K * 128, Cfg::B_BOX, B_SBO, B_LAYOUT);
              const uint32_t first = ((kb - kb0) | ks) != 0;
              if (NSPLIT == 1) {
                if (EB == 2) mma_f16(d_tmem, da, db, idesc, first); else mma_tf32(d_tmem, da, db, idesc, first);
              } else {
                const uint64_t da_lo = make_smem_desc(a0 + Cfg::A_TILE + ks * 32, 16, 1024);
                const uint64_t db_lo = make_smem_desc(b0 + Cfg::B_TILE + ks * Cfg::UMMA_K * 128, Cfg::B_BOX, B_SBO, B_LAYOUT);
                mma_tf32(d_tmem, da_lo, db, idesc, first);      // small terms first
                mma_tf32(d_tmem, da, db_lo, idesc, 1u);
                mma_tf32(d_tmem, da, db, idesc, 1u);
              }
            }
            mma_commit(&empty_bar[stage]);                      // smem stage reusable once these MMAs retire
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          mma_commit(&tmem_full[as]);                           // chunk accumulator complete -> epilogue
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5; TMEM lane quarter = warp % 4) =====================
    const int q = warp & 3;
    int ci = 0;
    const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      int tm, tn; tile_coord(t, tm, tn);
      const int row = tm * BM + q * 32 + lane;
      const bool row_ok = row < p.m;
      if constexpr (NSPLIT == 2) {
        // 3xTF32: promote every K chunk from TMEM into registers with round-to-nearest adds
        float accum[BN];
#pragma unroll
        for (int j = 0; j < BN; j++) accum[j] = 0.0f;
        for (int kb0 = 0; kb0 < kblocks; kb0 += chunk_blocks, ci++) {
          const int as = ci & 1;
          mbar_wait(&tmem_full[as], (ci >> 1) & 1);
          tc_fence_after();
#pragma unroll
          for (int c0 = 0; c0 < BN; c0 += 32) {
            uint32_t r[32];
            tmem_ld_32x32(tlane + (uint32_t)(as * BN + c0), r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j++) accum[c0 + j] = __fadd_rn(accum[c0 + j], __uint_as_float(r[j]));
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[as]);
        }
#pragma unroll
        for (int c0 = 0; c0 < BN; c0 += 32) {
          const int col0 = tn * BN + c0;
          if (row_ok && col0 < p.n) {
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; j++) v[j] = __fmul_rn(p.alpha, accum[c0 + j]);
            store_chunk<OUT, EPI>(p, v, row, col0);
          }
        }
      } else {
        // 16-bit routes: one accumulator stage per tile.  The chunk loop is NOT unrolled (the unrolled epilogue was
        // 230 KB of SASS and ran out of the instruction cache); the TMEM load of chunk i+1 is in flight while chunk i
        // is scaled, activated, packed and stored, and the TMEM stage is released as soon as the last chunk is in registers.
        const int as = ci & 1;
        mbar_wait(&tmem_full[as], (ci >> 1) & 1);
        tc_fence_after();
        ci++;
        const uint32_t tbase = tlane + (uint32_t)(as * BN);
        uint32_t r[32];
        tmem_ld_32x32(tbase, r);
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; j++) v[j] = __fmul_rn(p.alpha, __uint_as_float(r[j]));
          if (c0 + 32 < BN) {
            tmem_ld_32x32(tbase + (uint32_t)(c0 + 32), r);
          } else {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[as]);
          }
          const int col0 = tn * BN + c0;
          if (row_ok && col0 < p.n) store_chunk<OUT, EPI>(p, v, row, col0);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, Cfg::TMEM_COLS); }
}

// ---------------------------------------------------------------- CTA-pair (cta_group::2) variant, 16-bit operands
// One cluster of two CTAs (two SMs of a TPC) owns a 256 x 256 tile of C: tcgen05.mma.cta_group::2 with M = 256, N = 256.
// Each CTA stages its own 128 rows of A and HALF of the B tile (128 columns); the pair's tensor cores read both halves of B,
// so per SM and per 16-deep MMA step shared memory sees 8 KB of TMA writes + 8 KB of operand reads instead of 12 + 12 KB --
// the single-CTA kernel was shared-memory-bandwidth bound (tensor pipe 74 % active, ncu profiles/r01).
// Roles per CTA: warp 0 TMA producer (bytes counted on the LEADER's full barrier), warp 1 MMA issuer (leader CTA only; its
// tcgen05.commit is multicast to both CTAs' empty / tmem_full barriers), warps 2..5 epilogue on the CTA's own 128 TMEM lanes
// (the non-leader's epilogue warps release the accumulator stage on the leader's tmem_empty barrier).
struct TcPairCfg {
  static constexpr int BM = 128, BN = 256, BN_HALF = 128, BK = 64, UMMA_K = 16;
  static constexpr int A_TILE = BM * 128;                 // bytes: 128 rows x 128 B
  static constexpr int B_BOX = BK * 128;                  // 64 k-rows x 128 B (64 columns)
  static constexpr int B_TILE = (BN_HALF / 64) * B_BOX;   // this CTA's half of B
  static constexpr int STAGE = A_TILE + B_TILE;           // 32 KB
  static constexpr int STAGES = 6;
  static constexpr int ACC_STAGES = 2;
  static constexpr int TMEM_COLS = ACC_STAGES * BN;       // 512
  static constexpr int SMEM = STAGES * STAGE + 1024 + 256;
};

template <int OUT, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1)
gemm_tc_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, TcParams p) {
  using Cfg = TcPairCfg;
  constexpr int BM = Cfg::BM, BN = Cfg::BN, BK = Cfg::BK, STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + Cfg::ACC_STAGES;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + Cfg::ACC_STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();                 // 0 = leader
  const int num_tiles = p.tiles_m * p.tiles_n;             // tiles_m counts 256-row blocks
  const int kblocks = (p.k + BK - 1) / BK;
  const int first_tile = (int)cluster_id_x(), tile_step = (int)cluster_count_x();

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_a); prefetch_tmap(&map_b);
    for (int s = 0; s < STAGES; s++) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < Cfg::ACC_STAGES; s++) { mbar_init(&tmem_full[s], 1); mbar_init(&tmem_empty[s], 8); }
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc_pair(tmem_ptr, Cfg::TMEM_COLS); tmem_relinquish_pair(); }
  tc_fence_before();
  cluster_sync_all();                                      // barriers of both CTAs initialised before any remote arrive
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  auto tile_coord = [&](int t, int& tm, int& tn) {
    const int group = 4;                                   // 4 x 256 rows
    int per_group = group * p.tiles_n;
    int g = t / per_group, r = t % per_group;
    int rows = min(group, p.tiles_m - g * group);
    tm = g * group + r % rows;
    tn = r / rows;
  };

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = first_tile; t < num_tiles; t += tile_step) {
        int tm, tn; tile_coord(t, tm, tn);
        const int row0 = tm * 2 * BM + (int)rank * BM;
        const int col0 = tn * BN + (int)rank * Cfg::BN_HALF;
        for (int kb = 0; kb < kblocks; kb++) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* st = smem + stage * Cfg::STAGE;
          if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::STAGE);     // both CTAs' bytes land on the leader's barrier
          tma_load_2d_pair(st, &map_a, &full_bar[stage], kb * BK, row0);
          tma_load_2d_pair(st + Cfg::A_TILE, &map_b, &full_bar[stage], col0, kb * BK);
          tma_load_2d_pair(st + Cfg::A_TILE + Cfg::B_BOX, &map_b, &full_bar[stage], col0 + 64, kb * BK);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc(OUT == OUT_F16 ? 0u : 1u, /*a K-major*/ 0u, /*b MN-major*/ 1u, 2 * BM, BN);
      int stage = 0; uint32_t phase = 0;
      int ci = 0;
      for (int t = first_tile; t < num_tiles; t += tile_step, ci++) {
        const int as = ci & 1; const uint32_t aphase = (ci >> 1) & 1;
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < kblocks; kb++) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a0 = smem_u32(smem + stage * Cfg::STAGE);
          const uint32_t b0 = a0 + Cfg::A_TILE;
#pragma unroll
          for (int ks = 0; ks < BK / Cfg::UMMA_K; ks++) {
            const uint64_t da = make_smem_desc(a0 + ks * 32, 16, 1024);
            const uint64_t db = make_smem_desc(b0 + ks * Cfg::UMMA_K * 128, Cfg::B_BOX, 1024, 2u);
            mma_f16_pair(d_tmem, da, db, idesc, (uint32_t)((kb | ks) != 0));
          }
          mma_commit_pair(&empty_bar[stage]);                // the stage is reusable in BOTH CTAs once these MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        mma_commit_pair(&tmem_full[as]);                     // accumulator complete -> both CTAs' epilogues
      }
    }
  } else {
    // ===================== epilogue (warps 2..5 of both CTAs; TMEM lane quarter = warp % 4) =====================
    const int q = warp & 3;
    int ci = 0;
    const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int t = first_tile; t < num_tiles; t += tile_step, ci++) {
      int tm, tn; tile_coord(t, tm, tn);
      const int row = tm * 2 * BM + (int)rank * BM + q * 32 + lane;
      const bool row_ok = row < p.m;
      const int as = ci & 1;
      mbar_wait(&tmem_full[as], (ci >> 1) & 1);
      tc_fence_after();
      const uint32_t tbase = tlane + (uint32_t)(as * BN);
      uint32_t r[32];
      tmem_ld_32x32(tbase, r);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; j++) v[j] = __fmul_rn(p.alpha, __uint_as_float(r[j]));
        if (c0 + 32 < BN) {
          tmem_ld_32x32(tbase + (uint32_t)(c0 + 32), r);
        } else {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_leader(&tmem_empty[as]);
        }
        const int col0 = tn * BN + c0;
        if (row_ok && col0 < p.n) store_chunk<OUT, EPI>(p, v, row, col0);
      }
    }
  }
  tc_fence_before();
  cluster_sync_all();                                      // both CTAs done with TMEM and with each other's barriers
  if (warp == 1) { tc_fence_after(); tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS); }
}

// ---------------------------------------------------------------- host side
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;

static int get_encode() {
  static std::once_flag once;
  static int status = GUDA_OK;
  std::call_once(once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) status = GUDA_ERR_DEVICE;
    else g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  });
  if (status) return set_error(GUDA_ERR_DEVICE, "cuTensorMapEncodeTiled is not available from this driver");
  return GUDA_OK;
}

// 2-D row-major tensor [rows][cols] (cols contiguous, leading dimension ld elements), box = box_cols x box_rows, SWIZZLE_128B
static int make_map(CUtensorMap* map, CUtensorMapDataType dt, int eb, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                    uint32_t box_cols, uint32_t box_rows, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * (uint64_t)eb};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(GUDA_ERR_INVALID_ARG, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return GUDA_OK;
}

// plain (unswizzled) fp32 map for the SIMT pipeline of sgemm_pipe.cu: box rows are dense runs of box_cols floats
int tensor_map_2d_f32(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols, uint32_t box_rows) {
  int rc = get_encode();
  if (rc) return rc;
  return make_map(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, rows, cols, ld, box_cols, box_rows, CU_TENSOR_MAP_SWIZZLE_NONE);
}

template <int EB, int NSPLIT, int BN, int OUT, int EPI>
static int launch_tc(CUtensorMapDataType dt, int m, int n, int k, float alpha, const void* a0, const void* a1, long long lda,
                     const void* b0, const void* b1, long long ldb, float beta, void* c, long long ldc, const float* bias, cudaStream_t s) {
  using Cfg = TcCfg<EB, NSPLIT, BN>;
  int rc = get_encode();
  if (rc) return rc;
  CUtensorMap ma0, ma1, mb0, mb1;
  if ((rc = make_map(&ma0, dt, EB, a0, m, k, lda, Cfg::BK, Cfg::BM))) return rc;
  const CUtensorMapSwizzle bswz = EB == 4 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  if ((rc = make_map(&mb0, dt, EB, b0, k, n, ldb, Cfg::B_BOX_N, Cfg::BK, bswz))) return rc;
  ma1 = ma0; mb1 = mb0;
  if (NSPLIT == 2) {
    if ((rc = make_map(&ma1, dt, EB, a1, m, k, lda, Cfg::BK, Cfg::BM))) return rc;
    if ((rc = make_map(&mb1, dt, EB, b1, k, n, ldb, Cfg::B_BOX_N, Cfg::BK, bswz))) return rc;
  }
  TcParams p{m, n, k, alpha, beta, c, ldc, bias, EPI, (m + Cfg::BM - 1) / Cfg::BM, (n + BN - 1) / BN};
  auto kern = gemm_tc_kernel<EB, NSPLIT, BN, OUT, EPI>;
  static bool attr_set = false;
  if (!attr_set) {
    GUDA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    attr_set = true;
  }
  int grid = p.tiles_m * p.tiles_n;
  if (grid > sm_count()) grid = sm_count();
  kern<<<grid, 192, Cfg::SMEM, s>>>(ma0, ma1, mb0, mb1, p);
  return check_launch("gemm_tc");
}

template <int OUT, int EPI>
static int launch_tc_pair(CUtensorMapDataType dt, int m, int n, int k, float alpha, const void* a, long long lda, const void* b,
                          long long ldb, float beta, void* c, long long ldc, const float* bias, cudaStream_t s) {
  using Cfg = TcPairCfg;
  int rc = get_encode();
  if (rc) return rc;
  CUtensorMap ma, mb;
  if ((rc = make_map(&ma, dt, 2, a, m, k, lda, Cfg::BK, Cfg::BM))) return rc;
  if ((rc = make_map(&mb, dt, 2, b, k, n, ldb, 64, Cfg::BK))) return rc;
  TcParams p{m, n, k, alpha, beta, c, ldc, bias, EPI, (m + 2 * Cfg::BM - 1) / (2 * Cfg::BM), (n + Cfg::BN - 1) / Cfg::BN};
  auto kern = gemm_tc_pair_kernel<OUT, EPI>;
  static bool attr_set = false;
  if (!attr_set) {
    GUDA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    attr_set = true;
  }
  int clusters = p.tiles_m * p.tiles_n;
  if (clusters > sm_count() / 2) clusters = sm_count() / 2;
  kern<<<2 * clusters, 192, Cfg::SMEM, s>>>(ma, mb, p);      // __cluster_dims__(2,1,1): consecutive CTAs form a pair
  return check_launch("gemm_tc_pair");
}

static int g_tc_pair = 1;       // GUDA_TC_PAIR=0 selects the single-CTA kernel (A/B timing)

static bool tma_ok(const void* p, long long ld, int eb) { return ((uintptr_t)p % 16 == 0) && ((ld * eb) % 16 == 0); }

// 16-bit GEMM on tensor cores; -1 = not eligible (caller falls back to the exact route)
int gemm16_tc_try(int is_bf16, int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb, float beta,
                  uint16_t* c, int ldc, const float* bias, int epi, cudaStream_t s) {
  if (m < 128 || n < 128 || k < 64) return -1;                // tiny problems: the exact SIMT route is faster and bit-exact
  if (!tma_ok(a, lda, 2) || !tma_ok(b, ldb, 2)) return -1;
  static bool env_read = false;
  if (!env_read) { if (const char* e = getenv("GUDA_TC_PAIR")) g_tc_pair = atoi(e); env_read = true; }
  if (g_tc_pair && m >= 256 && n >= 256) {
    if (is_bf16) {
      const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
      if (epi == EPI_BIAS_RELU) return launch_tc_pair<OUT_BF16, EPI_BIAS_RELU>(dt, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, s);
      if (epi == EPI_BIAS_GELU) return launch_tc_pair<OUT_BF16, EPI_BIAS_GELU>(dt, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, s);
      return launch_tc_pair<OUT_BF16, EPI_NONE>(dt, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, s);
    }
    if (epi == EPI_NONE) return launch_tc_pair<OUT_F16, EPI_NONE>(CU_TENSOR_MAP_DATA_TYPE_FLOAT16, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, s);
  }
  if (is_bf16) {
    const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    if (epi == EPI_BIAS_RELU) return launch_tc<2, 1, 256, OUT_BF16, EPI_BIAS_RELU>(dt, m, n, k, alpha, a, a, lda, b, b, ldb, beta, c, ldc, bias, s);
    if (epi == EPI_BIAS_GELU) return launch_tc<2, 1, 256, OUT_BF16, EPI_BIAS_GELU>(dt, m, n, k, alpha, a, a, lda, b, b, ldb, beta, c, ldc, bias, s);
    return launch_tc<2, 1, 256, OUT_BF16, EPI_NONE>(dt, m, n, k, alpha, a, a, lda, b, b, ldb, beta, c, ldc, bias, s);
  }
  if (epi != EPI_NONE) return -1;
  return launch_tc<2, 1, 256, OUT_F16, EPI_NONE>(CU_TENSOR_MAP_DATA_TYPE_FLOAT16, m, n, k, alpha, a, a, lda, b, b, ldb, beta, c, ldc, bias, s);
}

// fp32 -> (hi, lo) planes for the 3xTF32 split: hi = x rounded to tf32 (10-bit mantissa, round-to-nearest-even on the
// 13 dropped bits), lo = x - hi (exact in fp32; the tensor core reads its top 10 mantissa bits)
__global__ void split_tf32_kernel(const float* __restrict__ src, long long lds, float* __restrict__ hi, float* __restrict__ lo,
                                  long long ldd, long long rows, long long cols) {
  const long long total = rows * cols;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long r = i / cols, c = i % cols;
    float x = src[r * lds + c];
    uint32_t u = __float_as_uint(x);
    float h = x;
    if ((u & 0x7F800000u) != 0x7F800000u) {            // finite: RNE to 10 mantissa bits
      uint32_t rb = u + 0x0FFFu + ((u >> 13) & 1u);
      h = __uint_as_float(rb & 0xFFFFE000u);
    }
    hi[r * ldd + c] = h;
    lo[r * ldd + c] = __fsub_rn(x, h);
  }
}


int sgemm_3xtf32_try(int ta, int tb, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb, float beta,
                     float* c, int ldc, cudaStream_t s) {
  if (ta || tb) return -1;
  if (m < 128 || n < 128 || k < 32) return -1;
  auto pad4 = [](long long x) { return (x + 3) & ~3ll; };
  const long long lda2 = pad4(k), ldb2 = pad4(n);
  PoolTemp ah, al, bh, bl;
  int rc;
  if ((rc = ah.alloc((size_t)m * lda2 * 4, s))) return rc;
  if ((rc = al.alloc((size_t)m * lda2 * 4, s))) return rc;
  if ((rc = bh.alloc((size_t)k * ldb2 * 4, s))) return rc;
  if ((rc = bl.alloc((size_t)k * ldb2 * 4, s))) return rc;
  const int blocks = sm_count() * 16;
  split_tf32_kernel<<<blocks, 256, 0, s>>>(a, lda, (float*)ah.p, (float*)al.p, lda2, m, k);
  split_tf32_kernel<<<blocks, 256, 0, s>>>(b, ldb, (float*)bh.p, (float*)bl.p, ldb2, k, n);
  if ((rc = check_launch("split_tf32"))) return rc;
  rc = launch_tc<4, 2, 128, OUT_F32, EPI_NONE>(CU_TENSOR_MAP_DATA_TYPE_FLOAT32, m, n, k, alpha, ah.p, al.p, lda2, bh.p, bl.p, ldb2, beta, c,
                                               ldc, nullptr, s);
  return rc;                                 // the planes return to the pool here, ordered after this stream's work
}

}  // namespace guda
