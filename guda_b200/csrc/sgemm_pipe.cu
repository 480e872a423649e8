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
#include "sgemm_pipe.cuh"
#include <cstdlib>
#include <mutex>

namespace guda {

int tensor_map_2d_f32(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols, uint32_t box_rows);

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
  pdl_wait();                // A may be the previous kernel's output, and the scratch it overwrites may still be read by it
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
  pdl_trigger();             // this CTA's part of At / flags is written: the pipeline kernel behind may start being scheduled
}

// ---------------------------------------------------------------- host side
// Configurations.  1 is the production choice; 2-4 stay selectable (guda_set_sgemm_pipe) as the A/B evidence DESIGN.md section 4 cites:
// the FFMA2 walk order decides how many operand reads the reuse cache absorbs (tools/explore_pipe.py), and that -- not shared
// memory, not the TMA ring -- is what bounds this kernel on sm_100a.
//                    TM  TN  LY  LX  WM WN  BK ST MINB ORDER DRY PAIR ACC64
using Pipe1  = PipeCfg< 8,  8,  2, 16,  4, 1, 32, 3, 2, 1, false, 0, false>;   //  64 x 128, 128 threads, 8 x 8 micro-tile, 32-deep stages: 67.0 TFLOP/s at 8192^3
using Pipe2  = PipeCfg< 8,  8,  2, 16,  4, 1, 32, 3, 2, 0, false, 1, false>;   //  the same tile, scalar-held walk order: 61.9
using Pipe3  = PipeCfg<16,  8,  2, 16,  4, 1, 16, 4, 2, 1, false, 1, false>;   // 128 x 128, 16 x 8 micro-tile: 65.4
using Pipe4  = PipeCfg< 8,  8,  2, 16,  4, 1, 32, 3, 2, 1, true,  0, false>;   //  Pipe1 with register-resident fragments (FFMA2-only timing experiment; wrong results)
constexpr int PIPE_NCFG = 4;
constexpr int PIPE_DEFAULT = 1;

static int g_pipe_cfg = [] { const char* e = getenv("GUDA_SGEMM_PIPE"); return e ? atoi(e) : -1; }();   // -1 auto, 0 off, 1.. forced config
static int g_pipe_lag = [] { const char* e = getenv("GUDA_SGEMM_PIPE_LAG"); return e ? atoi(e) : 1; }();
// L2 blocking of large calls.  A tile group is `rows` rows of C sweeping N: its A panel (rows x kc) stays in L2
// for the whole sweep, each B column tile (kc x 128) is shared by the group's rows/64 CTAs and re-read once per group.  The ~592
// resident CTAs sit at unrelated k positions, so the working set is the WHOLE resident patch: kc * (4*rows + 592*64*512/rows) bytes
// -- 144 MB at kc = 8192, rows = 2048 (round 1's depth; it thrashes the 126 MB L2: 2.74 TB of DRAM traffic per 65536^3 step) and
// 72 MB at kc = 4096, which fits: 1.31 TB per step at the same speed (profiles/r02/l2_blocking_sweep_32768.log: rows = 4096 falls
// off the cliff again).  Traffic model: 4*N^2*K/rows (B re-reads) + 8*N^2*K/kc (the C round trip of a K-chunked launch).
static int g_pipe_group_rows = [] { const char* e = getenv("GUDA_PIPE_GROUP_ROWS"); return e ? atoi(e) : 2048; }();
static int g_pipe_kc = [] { const char* e = getenv("GUDA_PIPE_KC"); return e ? atoi(e) : 4096; }();

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
  cudaError_t le = launch_pdl(4, kern, dim3((unsigned)tiles), dim3(Cfg::THREADS), Cfg::SMEM, s, ma, mb, p, flags, flag_ld, lag, g_pipe_group_rows, sp);
  if (le != cudaSuccess) return set_error(GUDA_ERR_EXECUTION, "sgemm_pipe launch failed: %s", cudaGetErrorString(le));
  return check_launch("sgemm_pipe");
}

template <class Cfg>
static bool pipe_fits(int m, int n, int k) { return m % Cfg::BM == 0 && n % Cfg::BN == 0 && k % Cfg::BK == 0; }

template <int EPI, bool SCAT>
static int launch_pipe(int cfg, const GemmParams<float>& p, const float* at, long long ldat, const unsigned char* flags, int flag_ld,
                       const ScatterParams<float, SCAT>& sp, cudaStream_t s) {
  switch (cfg) {
#define GUDA_PIPE_CASE(ID, CFG) case ID: return launch_pipe_cfg<CFG, EPI, SCAT>(p, at, ldat, flags, flag_ld, sp, s);
    GUDA_PIPE_CASE(1, Pipe1) GUDA_PIPE_CASE(2, Pipe2) GUDA_PIPE_CASE(3, Pipe3) GUDA_PIPE_CASE(4, Pipe4)
#undef GUDA_PIPE_CASE
  }
  return set_error(GUDA_ERR_INVALID_ARG, "sgemm_pipe: unknown configuration %d", cfg);
}

static bool pipe_cfg_fits(int cfg, int m, int n, int k) {
  switch (cfg) {
#define GUDA_PIPE_FIT(ID) case ID: return pipe_fits<Pipe##ID>(m, n, k);
    GUDA_PIPE_FIT(1) GUDA_PIPE_FIT(2) GUDA_PIPE_FIT(3) GUDA_PIPE_FIT(4)
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
  if (cfg < 0) {
    // automatic: the pipeline pays for its pre-pass and ring fill once the tile grid covers most of the machine
    cfg = PIPE_DEFAULT;
    if ((long long)(m / Pipe1::BM) * (n / Pipe1::BN) < 96 || k < 256) return -1;
  }
  if (!pipe_cfg_fits(cfg, m, n, k)) return -1;
  auto al16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  if (!al16(a) || !al16(b) || (c && !al16(c)) || lda % 4 || ldb % 4 || ldc % 4) return -1;
  if (scatter) {
    if (scatter->ld % 4 || scatter->off % 4) return -1;
    for (int i = 0; i < scatter->n; i++) if (!al16(scatter->dst[i])) return -1;
  }
  const int KC = g_pipe_kc;
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
    cudaError_t le = ta ? launch_pdl(2, sgemm_prepass_kernel<true>, pg, dim3(256), 0, s, a + (long long)k0 * lda, (long long)lda, m, kc, alpha, (float*)at.p,
                                     (long long)m, (unsigned char*)fl.p, flag_ld)
                        : launch_pdl(2, sgemm_prepass_kernel<false>, pg, dim3(256), 0, s, a + k0, (long long)lda, m, kc, alpha, (float*)at.p, (long long)m,
                                     (unsigned char*)fl.p, flag_ld);
    if (le != cudaSuccess) return set_error(GUDA_ERR_EXECUTION, "sgemm_prepass launch failed: %s", cudaGetErrorString(le));
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
