// level2.cu -- SGEMV (N / T) and SGER: HBM-streaming kernels whose arithmetic follows the reference chain for chain.
//
// Replaces compute/level2float32.go:244-308 (Sgemv) and :24-65 (Sger) with the package-f32 kernels underneath:
// compute/asm/f32/gemv.go:12-45 (GemvN), :52-91 (GemvT), compute/asm/f32/ge_amd64.s (Ger).
//
// Arithmetic contract (bit-exact against the oracle in every form):
//   GemvN, incX == incY == 1 : y_i = fl(y_i*beta) + fl(alpha * dot_i) (unfused), dot_i = DotUnitary(row_i, x) in the AVX2 order of
//       dotunitary_avx2_amd64.s: 32 FMA lanes over the 32-blocked body, (acc0+acc1)+(acc2+acc3), 16- and 8-wide FMA tails,
//       hi+lo fold, unfused 4-wide tail, (x0+x2),(x1+x3), h1+h0, scalar tail.  That order IS a warp: lane t owns the elements
//       i = t (mod 32), and the folds are shuffles by 8, 16, 4, 2, 1 -- so one warp per row reproduces the bits with fully
//       coalesced 128-byte row reads.
//   GemvN, strided           : dot_i = DotInc (dotinc_amd64.s): two unfused scalar accumulators (even/odd of each group of 4).
//   GemvT (all increments)   : y_j = beta == 0 ? 0 : fl(y_j*beta); then for i = 0..m-1: y_j = fl(fl(fl(alpha*x_i) * a_ij) + y_j)
//       (axpyunitaryto_amd64.s / axpyinc_amd64.s: MULPS + ADDPS).  A sequential chain per column: one lane per column, the rows
//       streamed through a 16-deep cp.async ring of 32x32 tiles so that a single chain still keeps ~60 KB in flight.
//   Ger                      : a_ij = fl(fl(fl(alpha*x_i) * y_j) + a_ij) (ge_amd64.s: MULPS + ADDPS in every block shape).
//
// Roofline: all three are HBM-bound.  Algorithmic bytes: GEMV 4*m*n (+ vectors), GER 8*m*n (read + write of A).
#include "common.cuh"
#include "tc_ptx.cuh"

namespace guda {
using tc::fence_barrier_init;
using tc::mbar_arrive;
using tc::mbar_init;
using tc::mbar_wait;
using tc::smem_u32;

__device__ __forceinline__ float ldg_stream_f32(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// ---------------------------------------------------------------------------------------------------- GemvN, unit increments
// One warp per row, U independent 128-byte row loads in flight per lane.  (Sharing the x loads between 2 or 4 rows of a warp was
// measured slower at every shape on B200 -- x stays in L1 -- so there is one instantiation.)
template <int U>
__global__ void __launch_bounds__(256) sgemv_n_unit_kernel(int m, int n, float alpha, const float* __restrict__ a, long long lda,
                                                           const float* __restrict__ x, float beta, float* __restrict__ y) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= m) return;
  const float* ap = a + row * lda;
  float acc = 0.0f;
  const int nb = n >> 5;
  int it = 0;
  for (; it + U <= nb; it += U) {
    float xv[U], av[U];
#pragma unroll
    for (int u = 0; u < U; u++) xv[u] = __ldg(x + (size_t)(it + u) * 32 + lane);
#pragma unroll
    for (int u = 0; u < U; u++) av[u] = ldg_stream_f32(ap + (size_t)(it + u) * 32 + lane);
#pragma unroll
    for (int u = 0; u < U; u++) acc = __fmaf_rn(av[u], xv[u], acc);
  }
  for (; it < nb; it++) acc = __fmaf_rn(ldg_stream_f32(ap + (size_t)it * 32 + lane), __ldg(x + (size_t)it * 32 + lane), acc);
  // lane t = 8*a + l holds acc[a][l] of the AVX2 kernel
  float v = __fadd_rn(acc, __shfl_down_sync(0xffffffffu, acc, 8));          // lanes 0-7: acc0+acc1, lanes 16-23: acc2+acc3
  float w = __fadd_rn(v, __shfl_down_sync(0xffffffffu, v, 16));             // lanes 0-7: (acc0+acc1)+(acc2+acc3)
  int idx = nb << 5, t = n & 31;
  if (t >= 16) {
    if (lane < 8) {
      w = __fmaf_rn(ap[idx + lane], x[idx + lane], w);
      w = __fmaf_rn(ap[idx + 8 + lane], x[idx + 8 + lane], w);
    }
    idx += 16; t -= 16;
  }
  if (t >= 8) {
    if (lane < 8) w = __fmaf_rn(ap[idx + lane], x[idx + lane], w);
    idx += 8; t -= 8;
  }
  float q = __fadd_rn(__shfl_down_sync(0xffffffffu, w, 4), w);              // lanes 0-3: hi + lo
  if (t >= 4) {
    if (lane < 4) q = __fadd_rn(q, __fmul_rn(ap[idx + lane], x[idx + lane]));
    idx += 4; t -= 4;
  }
  float h = __fadd_rn(q, __shfl_down_sync(0xffffffffu, q, 2));              // lane 0: x0+x2, lane 1: x1+x3
  float s = __fadd_rn(__shfl_down_sync(0xffffffffu, h, 1), h);              // lane 0: h1 + h0
  if (lane == 0) {
    for (; t > 0; t--, idx++) s = __fadd_rn(s, __fmul_rn(ap[idx], x[idx]));
    const float ad = __fmul_rn(alpha, s);
    float* yp = y + row;
    *yp = beta == 0.0f ? ad : __fadd_rn(__fmul_rn(*yp, beta), ad);
  }
}

// ---------------------------------------------------------------------------------------------------- GemvN, strided
// DotInc(x, row, n, incX, 1, kx, 0): sum takes elements 0 and 2 of each group of four, psum 1 and 3, all unfused.
__global__ void __launch_bounds__(128) sgemv_n_inc_kernel(int m, int n, float alpha, const float* __restrict__ a, long long lda,
                                                          const float* __restrict__ x, long long incx, long long kx, float beta,
                                                          float* __restrict__ y, long long incy, long long ky) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const float* row = a + i * lda;
  const float* px = x + kx;
  float sum = 0.0f;
  int j = 0;
  if (n >> 2) {
    float psum = 0.0f;
    for (; j + 4 <= n; j += 4) {
      const float p0 = __fmul_rn(px[0], row[j]), p1 = __fmul_rn(px[incx], row[j + 1]);
      const float p2 = __fmul_rn(px[2 * incx], row[j + 2]), p3 = __fmul_rn(px[3 * incx], row[j + 3]);
      sum = __fadd_rn(sum, p0); psum = __fadd_rn(psum, p1); sum = __fadd_rn(sum, p2); psum = __fadd_rn(psum, p3);
      px += 4 * incx;
    }
    sum = __fadd_rn(sum, psum);
  }
  for (; j < n; j++, px += incx) sum = __fadd_rn(sum, __fmul_rn(px[0], row[j]));
  const float ad = __fmul_rn(alpha, sum);
  float* yp = y + ky + i * incy;
  *yp = beta == 0.0f ? ad : __fadd_rn(__fmul_rn(*yp, beta), ad);
}

// ---------------------------------------------------------------------------------------------------- GemvT
// Two warps per CTA, specialised: a consumer whose lanes each own C adjacent columns (a 128*C-byte strip) and run the chains, and a
// producer that keeps the ring full (cp.async completion -> full[] mbarrier; the consumer releases stages through empty[]).
// A lone warp issues one instruction per ~2.5 cycles, so taking the address arithmetic and the copies off the chain warp roughly
// doubles the per-strip rate -- which is what bounds GemvT whenever n gives fewer strips than the GPU has schedulers.
// Rows arrive through a STAGES-deep cp.async ring of 32-row tiles; the stage's 32 x values ride the same ring and are read back
// four at a time as broadcast LDS.128.  With C = 2 the products are packed FMUL2
// (per-half IEEE; the adds stay scalar, see mul_add_x2), so the bits are still the unfused MULPS/ADDPS ones.
constexpr int GT_ROWS = 32;

__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_16_full(unsigned smem_addr, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_4(void* smem, const void* gmem, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// y0 = fl(fl(t*v0) + y0), y1 = fl(fl(t*v1) + y1): one packed multiply (FMUL2) and two scalar adds.  NOT mul.rn.f32x2 followed by
// add.rn.f32x2: ptxas 12.9 contracts that pair into a single FFMA2 (one rounding) even with explicit .rn and -fmad=false; the
// mixed packed/scalar form is left alone (checked in the SASS and by the bit-exact tests).
__device__ __forceinline__ void mul_add_x2(float& y0, float& y1, unsigned long long tt, float v0, float v1) {
  unsigned long long vv, pp;
  float p0, p1;
  asm("mov.b64 %0, {%1, %2};" : "=l"(vv) : "f"(v0), "f"(v1));
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(pp) : "l"(tt), "l"(vv));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(p0), "=f"(p1) : "l"(pp));
  y0 = __fadd_rn(p0, y0);
  y1 = __fadd_rn(p1, y1);
}

template <int C> struct GtVec;
template <> struct GtVec<1> { using type = float; };
template <> struct GtVec<2> { using type = float2; };

template <bool VEC, int C, int STAGES>
__global__ void __launch_bounds__(64) sgemv_t_kernel(int m, int n, float alpha, const float* __restrict__ a, long long lda,
                                                     const float* __restrict__ x, long long incx, long long kx, float beta,
                                                     float* __restrict__ y, long long incy, long long ky) {
  constexpr int COLS = 32 * C;
  constexpr int STAGE_FLOATS = GT_ROWS * COLS + GT_ROWS;           // the tile, then the stage's 32 x (later t) values
  extern __shared__ __align__(16) float gt_smem[];                 // [STAGES][STAGE_FLOATS], then full[STAGES], empty[STAGES]
  uint64_t* full = reinterpret_cast<uint64_t*>(gt_smem + (size_t)STAGES * STAGE_FLOATS);
  uint64_t* empty = full + STAGES;
  const int lane = threadIdx.x & 31;
  const int j0 = blockIdx.x * COLS;
  const int ntiles = (m + GT_ROWS - 1) / GT_ROWS;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&full[s], 32);                                     // one deferred arrive per producer lane
      mbar_init(&empty[s], 1);
    }
    fence_barrier_init();
  }
  __syncthreads();

  if (threadIdx.x >= 32) {
    // ---------------- producer warp: keeps the ring full; never touches the arithmetic
    const bool cols_full = j0 + COLS <= n;
    // 16-byte staging: 8*C lanes cover one tile row, one warp instruction covers RPI = 4/C rows, 8*C instructions a tile
    constexpr int LPR = 8 * C, RPI = 32 / LPR, NQ = GT_ROWS / RPI;
    const int vr = lane / LPR, vc4 = (lane % LPR) * 4;
    const float* vsrc = a + (long long)vr * lda + j0 + vc4;        // advanced by 32 rows per tile
    const float* xsrc = x + kx + (long long)lane * incx;
    const unsigned smem_base = smem_u32(gt_smem);
    const long long qstep = (long long)RPI * lda;
#pragma unroll 1
    for (int tile = 0; tile < ntiles; tile++) {
      const int st = tile % STAGES, use = tile / STAGES;
      if (use > 0) mbar_wait(&empty[st], (use - 1) & 1);
      float* dst = gt_smem + (size_t)st * STAGE_FLOATS;
      const int i0 = tile * GT_ROWS;
      {
        const bool ok = i0 + lane < m;
        cp_async_4(dst + GT_ROWS * COLS + lane, ok ? xsrc : x, ok ? 4 : 0);
        xsrc += GT_ROWS * incx;
      }
      if constexpr (VEC) {
        if (cols_full && i0 + GT_ROWS <= m) {
          unsigned d = smem_base + (unsigned)((st * STAGE_FLOATS + vr * COLS + vc4) * 4);
          const float* p = vsrc;
#pragma unroll
          for (int q = 0; q < NQ; q++) {
            cp_async_16_full(d, p);
            d += RPI * COLS * 4;
            p += qstep;
          }
        } else {
#pragma unroll 4
          for (int q = 0; q < NQ; q++) {
            const int r = q * RPI + vr;
            const int rem = n - (j0 + vc4);
            const bool ok = (i0 + r < m) && rem > 0;
            cp_async_16(dst + r * COLS + vc4, ok ? vsrc + (long long)q * qstep : a, ok ? (rem >= 4 ? 16 : rem * 4) : 0);
          }
        }
        vsrc += (long long)GT_ROWS * lda;
      } else {
#pragma unroll 4
        for (int r = 0; r < GT_ROWS; r++)
#pragma unroll
          for (int cc = 0; cc < C; cc++) {
            const int col = lane + 32 * cc;
            const bool ok = (j0 + col < n) && (i0 + r < m);
            cp_async_4(dst + r * COLS + col, ok ? a + (long long)(i0 + r) * lda + j0 + col : a, ok ? 4 : 0);
          }
      }
      // arrive on full[st] once this lane's copies above have landed
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&full[st])) : "memory");
    }
    cp_async_wait_all();
    return;
  }

  // ---------------- consumer warp: one lane per C columns, the chains
  const int jl = j0 + C * lane;                                    // this lane's first column
  float yv[C];
#pragma unroll
  for (int c = 0; c < C; c++) {
    yv[c] = 0.0f;
    if (jl + c < n && beta != 0.0f) yv[c] = __fmul_rn(y[ky + (long long)(jl + c) * incy], beta);
  }
  using V = typename GtVec<C>::type;

  // A tile is pulled into registers (32 row values per column + the 32 x values as 8 broadcast float4) one tile AHEAD of its
  // chain, so the shared-memory latency of tile k+1 hides under the 32 dependent adds of tile k; the stage is handed back to the
  // producer after that chain, when the loads have long landed.
  auto load = [&](int tile, V (&v)[GT_ROWS], float4 (&x4)[GT_ROWS / 4]) {
    const int st = tile % STAGES;
    mbar_wait(&full[st], (tile / STAGES) & 1);
    const float* stage = gt_smem + (size_t)st * STAGE_FLOATS;
    const V* src = reinterpret_cast<const V*>(stage) + lane;
#pragma unroll
    for (int r = 0; r < GT_ROWS; r++) v[r] = src[r * 32];
#pragma unroll
    for (int r4 = 0; r4 < GT_ROWS / 4; r4++) x4[r4] = reinterpret_cast<const float4*>(stage + GT_ROWS * COLS)[r4];
  };
  auto release = [&](int tile) {
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[tile % STAGES]);
  };
  auto step = [&](float xr, const V& v) {
    const float t = __fmul_rn(alpha, xr);                          // t_r = fl(alpha * x_r)
    if constexpr (C == 1) {
      yv[0] = __fadd_rn(__fmul_rn(t, v), yv[0]);
    } else {
      unsigned long long tt;
      asm("mov.b64 %0, {%1, %1};" : "=l"(tt) : "f"(t));
      mul_add_x2(yv[0], yv[1], tt, v.x, v.y);
    }
  };
  auto chain = [&](const V (&v)[GT_ROWS], const float4 (&x4)[GT_ROWS / 4]) {
#pragma unroll
    for (int r4 = 0; r4 < GT_ROWS / 4; r4++) {
      step(x4[r4].x, v[4 * r4 + 0]); step(x4[r4].y, v[4 * r4 + 1]); step(x4[r4].z, v[4 * r4 + 2]); step(x4[r4].w, v[4 * r4 + 3]);
    }
  };

  const int nfull = m / GT_ROWS;
  V va[GT_ROWS], vb[GT_ROWS];
  float4 xa[GT_ROWS / 4], xb[GT_ROWS / 4];
  int tile = 0;
  if (nfull > 0) load(0, va, xa);
#pragma unroll 1
  for (; tile + 2 <= nfull; tile += 2) {
    load(tile + 1, vb, xb);
    chain(va, xa);
    release(tile);
    if (tile + 2 < nfull) load(tile + 2, va, xa);
    chain(vb, xb);
    release(tile + 1);
  }
  if (tile < nfull) {
    chain(va, xa);
    release(tile);
    tile++;
  }
  if (tile < ntiles) {                                             // last, partial tile: rows beyond m must not enter the chain
    const int st = tile % STAGES;
    mbar_wait(&full[st], (tile / STAGES) & 1);
    const float* stage = gt_smem + (size_t)st * STAGE_FLOATS;
    const V* src = reinterpret_cast<const V*>(stage) + lane;
    const int rows = m - tile * GT_ROWS;
    for (int r = 0; r < rows; r++) step(stage[GT_ROWS * COLS + r], src[r * 32]);
  }
#pragma unroll
  for (int c = 0; c < C; c++)
    if (jl + c < n) y[ky + (long long)(jl + c) * incy] = yv[c];
}

// ---------------------------------------------------------------------------------------------------- Ger
// 128 columns x 32 rows per CTA, 128-bit accesses; each thread holds its four y values and walks 4 rows.
__global__ void __launch_bounds__(256) sger_vec_kernel(int m, int n, float alpha, const float* __restrict__ x, const float* __restrict__ y,
                                                       float* __restrict__ a, long long lda) {
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int j = (blockIdx.y * 32 + tx) * 4;
  const int i0 = blockIdx.x * 32 + ty;
  if (j >= n) return;
  const float4 yv = *reinterpret_cast<const float4*>(y + j);
  float4 av[4];
  float t[4];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int i = i0 + 8 * q;
    if (i < m) {
      av[q] = *reinterpret_cast<const float4*>(a + (long long)i * lda + j);
      t[q] = __fmul_rn(alpha, __ldg(x + i));
    }
  }
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int i = i0 + 8 * q;
    if (i < m) {
      float4 r;
      r.x = __fadd_rn(__fmul_rn(t[q], yv.x), av[q].x);
      r.y = __fadd_rn(__fmul_rn(t[q], yv.y), av[q].y);
      r.z = __fadd_rn(__fmul_rn(t[q], yv.z), av[q].z);
      r.w = __fadd_rn(__fmul_rn(t[q], yv.w), av[q].w);
      *reinterpret_cast<float4*>(a + (long long)i * lda + j) = r;
    }
  }
}

__global__ void __launch_bounds__(256) sger_scalar_kernel(int m, int n, float alpha, const float* __restrict__ x, long long incx, long long kx,
                                                          const float* __restrict__ y, long long incy, long long ky,
                                                          float* __restrict__ a, long long lda) {
  const int j = blockIdx.y * 32 + (threadIdx.x & 31);
  const int i0 = blockIdx.x * 32 + (threadIdx.x >> 5);
  if (j >= n) return;
  const float yv = y[ky + (long long)j * incy];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int i = i0 + 8 * q;
    if (i < m) {
      const float t = __fmul_rn(alpha, x[kx + (long long)i * incx]);
      float* p = a + (long long)i * lda + j;
      *p = __fadd_rn(__fmul_rn(t, yv), *p);
    }
  }
}

// ---------------------------------------------------------------------------------------------------- fp64: Dgemv / Dger
// compute/asm/f64/gemvN_avx2_amd64.s: per row four FMA lanes (lane l owns the elements j = l mod 4 of the 4-blocked body), folded
// (l0+l2)+(l1+l3), unfused remainder, y_i = fl(alpha*s) + (beta == 0 ? 0 : fl(y_i*beta)).  Four threads per row, eight rows per
// warp: each lane replays its AVX lane's chain, the folds are shuffles by 2 and 1.
__global__ void __launch_bounds__(256) dgemv_n_unit_kernel(int m, int n, double alpha, const double* __restrict__ a, long long lda,
                                                           const double* __restrict__ x, double beta, double* __restrict__ y, long long incy,
                                                           long long ky) {
  const long long row = (long long)blockIdx.x * 64 + (threadIdx.x >> 2);
  const int l = threadIdx.x & 3;
  const double* ap = a + (row < m ? row : m - 1) * lda;
  double acc = 0.0;
  const int n4 = n >> 2;
  int it = 0;
  for (; it + 16 <= n4; it += 16) {
    double av[16], xv[16];
#pragma unroll
    for (int u = 0; u < 16; u++) { av[u] = ap[4 * (it + u) + l]; xv[u] = x[4 * (it + u) + l]; }
#pragma unroll
    for (int u = 0; u < 16; u++) acc = __fma_rn(av[u], xv[u], acc);
  }
  for (; it < n4; it++) acc = __fma_rn(ap[4 * it + l], x[4 * it + l], acc);
  const double v = __dadd_rn(acc, __shfl_down_sync(0xffffffffu, acc, 2));      // l = 0: l0 + l2, l = 1: l1 + l3
  double s = __dadd_rn(v, __shfl_down_sync(0xffffffffu, v, 1));                // l = 0: (l0 + l2) + (l1 + l3)
  if (l == 0 && row < m) {
    for (int j = n4 << 2; j < n; j++) s = __dadd_rn(s, __dmul_rn(ap[j], x[j]));
    double* yp = y + ky + row * incy;
    const double ys = beta == 0.0 ? 0.0 : __dmul_rn(*yp, beta);
    *yp = __dadd_rn(__dmul_rn(alpha, s), ys);
  }
}

// The same chains fed through shared memory: m rows supply only 4m chains, far too few 8-byte loads in flight to cover HBM latency
// from registers (8192^2: 2.96 TB/s).  One warp per CTA owns 8 rows; all 32 lanes stream 8 x 64-column tiles (and the tile's 64 x
// values) with 16-byte cp.async into a DN_STAGES-deep ring, 512 contiguous bytes of one row per instruction, and lane (r, l) then
// runs its chain out of the stage.  Rows are padded by 32 bytes so the 8 rows' 32-byte groups land in distinct banks (two
// wavefronts per 256-byte warp read, the minimum).  Needs 16-byte aligned rows and x (else dgemv_n_unit_kernel above).
constexpr int DN_STAGES = 6, DN_COLS = 64, DN_ROW = DN_COLS + 4, DN_STAGE_DOUBLES = 8 * DN_ROW + DN_COLS;

__global__ void __launch_bounds__(32) dgemv_n_ring_kernel(int m, int n, double alpha, const double* __restrict__ a, long long lda,
                                                          const double* __restrict__ x, double beta, double* __restrict__ y, long long incy,
                                                          long long ky) {
  __shared__ __align__(16) double ring[DN_STAGES * DN_STAGE_DOUBLES];
  const int lane = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * 8;
  const int n4 = n >> 2, nbody = n4 << 2;
  const int ntiles = (nbody + DN_COLS - 1) / DN_COLS;
  // producer view: lane covers the 16-byte chunk `lane` of a row's 512-byte tile segment, one instruction per row
  const double* src[8];
#pragma unroll
  for (int r = 0; r < 8; r++) src[r] = a + (row0 + r < m ? row0 + r : (long long)m - 1) * lda + 2 * lane;
  const unsigned ring_base = smem_u32(ring);
  auto issue = [&](int tile) {
    if (tile < ntiles) {
      const int st = tile % DN_STAGES;
      const int c0 = tile * DN_COLS;
      const unsigned d = ring_base + (unsigned)((st * DN_STAGE_DOUBLES + 2 * lane) * 8);
      if (c0 + DN_COLS <= nbody) {
#pragma unroll
        for (int r = 0; r < 8; r++) cp_async_16_full(d + r * DN_ROW * 8, src[r] + c0);
        if (lane < DN_COLS / 2) cp_async_16_full(d + 8 * DN_ROW * 8, x + c0 + 2 * lane);
      } else if (c0 + 2 * lane < nbody) {                               // nbody is a multiple of 4: chunks are whole or absent
#pragma unroll
        for (int r = 0; r < 8; r++) cp_async_16_full(d + r * DN_ROW * 8, src[r] + c0);
        cp_async_16_full(d + 8 * DN_ROW * 8, x + c0 + 2 * lane);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int t = 0; t < DN_STAGES - 1; t++) issue(t);

  const int r = lane >> 2, l = lane & 3;
  double acc = 0.0;
#pragma unroll 1
  for (int tile = 0; tile < ntiles; tile++) {
    asm volatile("cp.async.wait_group %0;" ::"n"(DN_STAGES - 2) : "memory");
    __syncwarp();                                                       // tile landed for every lane; all lanes are past tile - 1,
    issue(tile + DN_STAGES - 1);                                        // whose stage this refills
    const double* stage = ring + (size_t)(tile % DN_STAGES) * DN_STAGE_DOUBLES;
    const double* ar = stage + r * DN_ROW + l;
    const double* xr = stage + 8 * DN_ROW + l;
    const int nb = min(DN_COLS / 4, n4 - tile * (DN_COLS / 4));
    if (nb == DN_COLS / 4) {
      double av[DN_COLS / 4], xv[DN_COLS / 4];
#pragma unroll
      for (int u = 0; u < DN_COLS / 4; u++) { av[u] = ar[4 * u]; xv[u] = xr[4 * u]; }
#pragma unroll
      for (int u = 0; u < DN_COLS / 4; u++) acc = __fma_rn(av[u], xv[u], acc);
    } else {
      for (int u = 0; u < nb; u++) acc = __fma_rn(ar[4 * u], xr[4 * u], acc);
    }
  }
  cp_async_wait_all();
  const double v = __dadd_rn(acc, __shfl_down_sync(0xffffffffu, acc, 2));
  double s = __dadd_rn(v, __shfl_down_sync(0xffffffffu, v, 1));
  const long long row = row0 + r;
  if (l == 0 && row < m) {
    const double* ap = a + row * lda;
    for (int j = nbody; j < n; j++) s = __dadd_rn(s, __dmul_rn(ap[j], x[j]));
    double* yp = y + ky + row * incy;
    const double ys = beta == 0.0 ? 0.0 : __dmul_rn(*yp, beta);
    *yp = __dadd_rn(__dmul_rn(alpha, s), ys);
  }
}

// incX != 1: one unfused scalar chain per row, then the same fold with empty upper lanes ((s + 0) + (0 + 0))
__global__ void __launch_bounds__(128) dgemv_n_inc_kernel(int m, int n, double alpha, const double* __restrict__ a, long long lda,
                                                          const double* __restrict__ x, long long incx, long long kx, double beta,
                                                          double* __restrict__ y, long long incy, long long ky) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const double* row = a + i * lda;
  const double* px = x + kx;
  double s = 0.0;
  for (int j = 0; j < n; j++, px += incx) s = __dadd_rn(s, __dmul_rn(row[j], px[0]));
  s = __dadd_rn(__dadd_rn(s, 0.0), 0.0);
  double* yp = y + ky + i * incy;
  const double ys = beta == 0.0 ? 0.0 : __dmul_rn(*yp, beta);
  *yp = __dadd_rn(__dmul_rn(alpha, s), ys);
}

// compute/asm/f64/gemvT_amd64.s: per column the sequential unfused chain y_j += fl(fl(alpha*x_i) * a_ij) in row order.  One thread
// per column (coalesced across columns), 32 rows of loads in flight.
__global__ void __launch_bounds__(64) dgemv_t_kernel(int m, int n, double alpha, const double* __restrict__ a, long long lda,
                                                      const double* __restrict__ x, long long incx, long long kx, double beta,
                                                      double* __restrict__ y, long long incy, long long ky) {
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  double* yp = y + ky + j * incy;
  double yv = beta == 0.0 ? 0.0 : (beta == 1.0 ? *yp : __dmul_rn(*yp, beta));
  const double* ap = a + j;
  const double* px = x + kx;
  int i = 0;
  for (; i + 32 <= m; i += 32) {
    double av[32];
#pragma unroll
    for (int u = 0; u < 32; u++) av[u] = ap[(long long)(i + u) * lda];
#pragma unroll
    for (int u = 0; u < 32; u++) yv = __dadd_rn(yv, __dmul_rn(__dmul_rn(alpha, px[(long long)(i + u) * incx]), av[u]));
  }
  for (; i < m; i++) yv = __dadd_rn(yv, __dmul_rn(__dmul_rn(alpha, px[(long long)i * incx]), ap[(long long)i * lda]));
  *yp = yv;
}


// The fp64 GemvT on the same two-warp ring as sgemv_t_kernel (one column per lane, 256-byte strips, 8 stages of 32 rows): the
// producer stages 16-byte chunks (two doubles) and the tile's 32 x values, the consumer holds tile k+1 in registers while it runs
// tile k's 32 dependent adds.  Needs 16-byte aligned rows (a % 16 == 0, lda even); dgemv_t_kernel above is the fallback.
constexpr int DT_STAGES = 8;
constexpr int DT_STAGE_DOUBLES = GT_ROWS * 32 + GT_ROWS;

__device__ __forceinline__ void cp_async_8(void* smem, const void* gmem, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes) : "memory");
}

__global__ void __launch_bounds__(64) dgemv_t_ring_kernel(int m, int n, double alpha, const double* __restrict__ a, long long lda,
                                                          const double* __restrict__ x, long long incx, long long kx, double beta,
                                                          double* __restrict__ y, long long incy, long long ky) {
  extern __shared__ __align__(16) double dt_smem[];                // [DT_STAGES][DT_STAGE_DOUBLES], then full[], empty[]
  uint64_t* full = reinterpret_cast<uint64_t*>(dt_smem + (size_t)DT_STAGES * DT_STAGE_DOUBLES);
  uint64_t* empty = full + DT_STAGES;
  const int lane = threadIdx.x & 31;
  const int j0 = blockIdx.x * 32;
  const int ntiles = (m + GT_ROWS - 1) / GT_ROWS;
  if (threadIdx.x == 0) {
    for (int s = 0; s < DT_STAGES; s++) {
      mbar_init(&full[s], 32);
      mbar_init(&empty[s], 1);
    }
    fence_barrier_init();
  }
  __syncthreads();

  if (threadIdx.x >= 32) {
    // ---------------- producer warp: 16 lanes cover one 256-byte tile row, one instruction two rows, 16 instructions a tile
    const bool cols_full = j0 + 32 <= n;
    const int vr = lane >> 4, vc = (lane & 15) * 2;
    const double* vsrc = a + (long long)vr * lda + j0 + vc;
    const double* xsrc = x + kx + (long long)lane * incx;
    const unsigned smem_base = smem_u32(dt_smem);
    const long long qstep = 2 * lda;
#pragma unroll 1
    for (int tile = 0; tile < ntiles; tile++) {
      const int st = tile % DT_STAGES, use = tile / DT_STAGES;
      if (use > 0) mbar_wait(&empty[st], (use - 1) & 1);
      double* dst = dt_smem + (size_t)st * DT_STAGE_DOUBLES;
      const int i0 = tile * GT_ROWS;
      {
        const bool ok = i0 + lane < m;
        cp_async_8(dst + GT_ROWS * 32 + lane, ok ? xsrc : x, ok ? 8 : 0);
        xsrc += GT_ROWS * incx;
      }
      if (cols_full && i0 + GT_ROWS <= m) {
        unsigned d = smem_base + (unsigned)((st * DT_STAGE_DOUBLES + vr * 32 + vc) * 8);
        const double* p = vsrc;
#pragma unroll
        for (int q = 0; q < 16; q++) {
          cp_async_16_full(d, p);
          d += 2 * 32 * 8;
          p += qstep;
        }
      } else {
#pragma unroll 4
        for (int q = 0; q < 16; q++) {
          const int r = q * 2 + vr;
          const int rem = n - (j0 + vc);
          const bool ok = (i0 + r < m) && rem > 0;
          cp_async_16(dst + r * 32 + vc, ok ? vsrc + (long long)q * qstep : a, ok ? (rem >= 2 ? 16 : 8) : 0);
        }
      }
      vsrc += (long long)GT_ROWS * lda;
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&full[st])) : "memory");
    }
    cp_async_wait_all();
    return;
  }

  // ---------------- consumer warp
  const int j = j0 + lane;
  double* yp = y + ky + (long long)j * incy;
  double yv = 0.0;
  if (j < n && beta != 0.0) yv = beta == 1.0 ? *yp : __dmul_rn(*yp, beta);
  auto load = [&](int tile, double (&v)[GT_ROWS]) {
    const int st = tile % DT_STAGES;
    mbar_wait(&full[st], (tile / DT_STAGES) & 1);
    const double* stage = dt_smem + (size_t)st * DT_STAGE_DOUBLES;
#pragma unroll
    for (int r = 0; r < GT_ROWS; r++) v[r] = stage[r * 32 + lane];
  };
  auto release = [&](int tile) {
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[tile % DT_STAGES]);
  };
  auto step = [&](double xr, double v) { yv = __dadd_rn(yv, __dmul_rn(__dmul_rn(alpha, xr), v)); };
  // the x values stay in the stage (the tile's 64 registers per buffer are all a double-precision chain can afford) and are read
  // as broadcast 128-bit loads while the chain runs; the stage is released after the chain either way
  auto chain = [&](int tile, const double (&v)[GT_ROWS]) {
    const double2* xs = reinterpret_cast<const double2*>(dt_smem + (size_t)(tile % DT_STAGES) * DT_STAGE_DOUBLES + GT_ROWS * 32);
#pragma unroll
    for (int r2 = 0; r2 < GT_ROWS / 2; r2++) {
      const double2 x2 = xs[r2];
      step(x2.x, v[2 * r2]); step(x2.y, v[2 * r2 + 1]);
    }
  };
  const int nfull = m / GT_ROWS;
  double va[GT_ROWS], vb[GT_ROWS];
  int tile = 0;
  if (nfull > 0) load(0, va);
#pragma unroll 1
  for (; tile + 2 <= nfull; tile += 2) {
    load(tile + 1, vb);
    chain(tile, va);
    release(tile);
    if (tile + 2 < nfull) load(tile + 2, va);
    chain(tile + 1, vb);
    release(tile + 1);
  }
  if (tile < nfull) {
    chain(tile, va);
    release(tile);
    tile++;
  }
  if (tile < ntiles) {
    const int st = tile % DT_STAGES;
    mbar_wait(&full[st], (tile / DT_STAGES) & 1);
    const double* stage = dt_smem + (size_t)st * DT_STAGE_DOUBLES;
    const int rows = m - tile * GT_ROWS;
    for (int r = 0; r < rows; r++) step(stage[GT_ROWS * 32 + r], stage[r * 32 + lane]);
  }
  if (j < n) *yp = yv;
}

// compute/asm/f64/ger_amd64.s: a_ij = fl(fl(alpha*x_i) * y_j) + a_ij (MULPD + ADDPD in every block shape)
__global__ void __launch_bounds__(256) dger_kernel(int m, int n, double alpha, const double* __restrict__ x, long long incx, long long kx,
                                                   const double* __restrict__ y, long long incy, long long ky, double* __restrict__ a,
                                                   long long lda) {
  const int j = blockIdx.y * 32 + (threadIdx.x & 31);
  const int i0 = blockIdx.x * 32 + (threadIdx.x >> 5);
  if (j >= n) return;
  const double yv = y[ky + (long long)j * incy];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int i = i0 + 8 * q;
    if (i < m) {
      const double t = __dmul_rn(alpha, x[kx + (long long)i * incx]);
      double* p = a + (long long)i * lda + j;
      *p = __dadd_rn(__dmul_rn(t, yv), *p);
    }
  }
}

// Columns per lane for GemvT.  The chains are sequential in m, so the kernel's time is ~ m * (cycles per row of the consumer warp)
// until n supplies more strips than the GPU has schedulers (592); past that, two columns per lane (fewer instructions per byte)
// is what reaches the HBM roofline.  Measured on B200: C = 1 wins for n <= 8192, C = 2 from n = 16384 up.
static int gemvt_cols_per_lane(int n) { return n >= 16384 ? 2 : 1; }

}  // namespace guda

using namespace guda;
extern "C" {

int guda_sgemv(int trans, int m, int n, float alpha, const float* a, int lda, const float* x, int incx, float beta, float* y, int incy,
               guda_stream_t stream) {
  if (m < 0) GUDA_PANIC("blas: m < 0");
  if (n < 0) GUDA_PANIC("blas: n < 0");
  if (lda < (n > 1 ? n : 1)) GUDA_PANIC("blas: bad leading dimension of A");
  if (incx == 0) GUDA_PANIC("blas: zero x index increment");
  if (incy == 0) GUDA_PANIC("blas: zero y index increment");
  if (m == 0 || n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!a || !x || !y) return set_error(GUDA_ERR_INVALID_ARG, "Sgemv: null pointer");
  if (alpha == 0.0f && beta == 1.0f) return GUDA_OK;
  const int lenx = trans ? m : n, leny = trans ? n : m;
  if (alpha == 0.0f) return guda_sscal(leny, beta, y, incy > 0 ? incy : -incy, stream);      // level2float32.go:291-299
  cudaStream_t s = native(stream);
  const long long kx = incx < 0 ? -(long long)(lenx - 1) * incx : 0, ky = incy < 0 ? -(long long)(leny - 1) * incy : 0;
  if (!trans) {
    if (incx == 1 && incy == 1) {
      sgemv_n_unit_kernel<8><<<(unsigned)(((long long)m + 7) / 8), 256, 0, s>>>(m, n, alpha, a, lda, x, beta, y);
    } else {
      sgemv_n_inc_kernel<<<(unsigned)((m + 127) / 128), 128, 0, s>>>(m, n, alpha, a, lda, x, incx, kx, beta, y, incy, ky);
    }
    return check_launch("Sgemv");
  }
  const bool vec = (((uintptr_t)a & 15) == 0) && (lda % 4 == 0);
  const char* e = getenv("GUDA_GEMVT_COLS");            // read per call: tests and A/B runs flip it in-process
  int cpl = e ? atoi(e) : 0;
  if (cpl != 1 && cpl != 2) cpl = gemvt_cols_per_lane(n);
#define GT_LAUNCH(CPL, S)                                                                                                        \
  do {                                                                                                                           \
    const size_t smem = (size_t)(S) * ((GT_ROWS * 32 * (CPL) + GT_ROWS) * sizeof(float) + 16);                                          \
    static bool attr_done = false;                                                                                               \
    if (!attr_done) {                                                                                                            \
      GUDA_CUDA(cudaFuncSetAttribute(sgemv_t_kernel<true, CPL, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));      \
      GUDA_CUDA(cudaFuncSetAttribute(sgemv_t_kernel<false, CPL, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));     \
      attr_done = true;                                                                                                          \
    }                                                                                                                            \
    const unsigned grid = (unsigned)((n + 32 * (CPL) - 1) / (32 * (CPL)));                                                       \
    if (vec) sgemv_t_kernel<true, CPL, S><<<grid, 64, smem, s>>>(m, n, alpha, a, lda, x, incx, kx, beta, y, incy, ky);            \
    else sgemv_t_kernel<false, CPL, S><<<grid, 64, smem, s>>>(m, n, alpha, a, lda, x, incx, kx, beta, y, incy, ky);               \
  } while (0)
  if (cpl == 2) GT_LAUNCH(2, 6);
  else GT_LAUNCH(1, 8);
#undef GT_LAUNCH
  return check_launch("Sgemv");
}

int guda_dgemv(int trans, int m, int n, double alpha, const double* a, int lda, const double* x, int incx, double beta, double* y, int incy,
               guda_stream_t stream) {
  if (m < 0) GUDA_PANIC("blas: m < 0");
  if (n < 0) GUDA_PANIC("blas: n < 0");
  if (lda < (n > 1 ? n : 1)) GUDA_PANIC("blas: bad leading dimension of A");
  if (incx == 0) GUDA_PANIC("blas: zero x index increment");
  if (incy == 0) GUDA_PANIC("blas: zero y index increment");
  if (m == 0 || n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!a || !x || !y) return set_error(GUDA_ERR_INVALID_ARG, "Dgemv: null pointer");
  if (alpha == 0.0 && beta == 1.0) return GUDA_OK;
  const int lenx = trans ? m : n, leny = trans ? n : m;
  if (alpha == 0.0) return guda_dscal(leny, beta, y, incy > 0 ? incy : -incy, stream);      // compute/level2float64.go:279-287
  cudaStream_t s = native(stream);
  const long long kx = incx < 0 ? -(long long)(lenx - 1) * incx : 0, ky = incy < 0 ? -(long long)(leny - 1) * incy : 0;
  if (!trans) {
    if (incx == 1 && (((uintptr_t)a | (uintptr_t)x) & 15) == 0 && lda % 2 == 0 && n >= 4 * DN_COLS && !getenv("GUDA_DGEMVN_SIMPLE"))
      dgemv_n_ring_kernel<<<(unsigned)(((long long)m + 7) / 8), 32, 0, s>>>(m, n, alpha, a, lda, x, beta, y, incy, ky);
    else if (incx == 1) dgemv_n_unit_kernel<<<(unsigned)(((long long)m + 63) / 64), 256, 0, s>>>(m, n, alpha, a, lda, x, beta, y, incy, ky);
    else dgemv_n_inc_kernel<<<(unsigned)((m + 127) / 128), 128, 0, s>>>(m, n, alpha, a, lda, x, incx, kx, beta, y, incy, ky);
  } else if ((((uintptr_t)a & 15) == 0) && (lda % 2 == 0) && !getenv("GUDA_DGEMVT_SIMPLE")) {
    const size_t smem = (size_t)DT_STAGES * (DT_STAGE_DOUBLES * sizeof(double) + 16);
    static bool attr_done = false;
    if (!attr_done) {
      GUDA_CUDA(cudaFuncSetAttribute(dgemv_t_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_done = true;
    }
    dgemv_t_ring_kernel<<<(unsigned)((n + 31) / 32), 64, smem, s>>>(m, n, alpha, a, lda, x, incx, kx, beta, y, incy, ky);
  } else {
    dgemv_t_kernel<<<(unsigned)((n + 63) / 64), 64, 0, s>>>(m, n, alpha, a, lda, x, incx, kx, beta, y, incy, ky);
  }
  return check_launch("Dgemv");
}

int guda_dger(int m, int n, double alpha, const double* x, int incx, const double* y, int incy, double* a, int lda, guda_stream_t stream) {
  if (m < 0) GUDA_PANIC("blas: m < 0");
  if (n < 0) GUDA_PANIC("blas: n < 0");
  if (lda < (n > 1 ? n : 1)) GUDA_PANIC("blas: bad leading dimension of A");
  if (incx == 0) GUDA_PANIC("blas: zero x index increment");
  if (incy == 0) GUDA_PANIC("blas: zero y index increment");
  if (m == 0 || n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!a || !x || !y) return set_error(GUDA_ERR_INVALID_ARG, "Dger: null pointer");
  if (alpha == 0.0) return GUDA_OK;
  const long long kx = incx < 0 ? -(long long)(m - 1) * incx : 0, ky = incy < 0 ? -(long long)(n - 1) * incy : 0;
  dim3 grid((unsigned)((m + 31) / 32), (unsigned)((n + 31) / 32));
  if (grid.y > 65535u) return set_error(GUDA_ERR_INVALID_ARG, "Dger: n too large for one launch");
  dger_kernel<<<grid, 256, 0, native(stream)>>>(m, n, alpha, x, incx, kx, y, incy, ky, a, lda);
  return check_launch("Dger");
}

int guda_sger(int m, int n, float alpha, const float* x, int incx, const float* y, int incy, float* a, int lda, guda_stream_t stream) {
  if (m < 0) GUDA_PANIC("blas: m < 0");
  if (n < 0) GUDA_PANIC("blas: n < 0");
  if (lda < (n > 1 ? n : 1)) GUDA_PANIC("blas: bad leading dimension of A");
  if (incx == 0) GUDA_PANIC("blas: zero x index increment");
  if (incy == 0) GUDA_PANIC("blas: zero y index increment");
  if (m == 0 || n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!a || !x || !y) return set_error(GUDA_ERR_INVALID_ARG, "Sger: null pointer");
  if (alpha == 0.0f) return GUDA_OK;
  cudaStream_t s = native(stream);
  const bool vec = incx == 1 && incy == 1 && (n % 4 == 0) && (lda % 4 == 0) && (((uintptr_t)a | (uintptr_t)y) & 15) == 0;
  if (vec) {
    dim3 grid((unsigned)((m + 31) / 32), (unsigned)((n / 4 + 31) / 32));
    if (grid.y > 65535u) return set_error(GUDA_ERR_INVALID_ARG, "Sger: n too large for one launch");
    sger_vec_kernel<<<grid, 256, 0, s>>>(m, n, alpha, x, y, a, lda);
  } else {
    const long long kx = incx < 0 ? -(long long)(m - 1) * incx : 0, ky = incy < 0 ? -(long long)(n - 1) * incy : 0;
    dim3 grid((unsigned)((m + 31) / 32), (unsigned)((n + 31) / 32));
    if (grid.y > 65535u) return set_error(GUDA_ERR_INVALID_ARG, "Sger: n too large for one launch");
    sger_scalar_kernel<<<grid, 256, 0, s>>>(m, n, alpha, x, incx, kx, y, incy, ky, a, lda);
  }
  return check_launch("Sger");
}
}
