// conv.cu -- Conv2D (NCHW) as im2col -> exact-order SGEMM -> NCHW placement + bias.  Replaces conv.go:69-226.
//
// Same arithmetic as the reference: out[b][oc][p] = fl( chain_k fma(w[oc][k], col[k][b*P+p]) + bias[oc] ), the chain being
// Sgemm(NoTrans, NoTrans, OC, B*P, K, 1, kernel, K, col, B*P, 0, ...) of compute/sgemm.go with k = (c, kh, kw) ascending
// and the reference's zero test on the weights -- so for BatchSize == 1 (all the reference's tests) results are bit-identical
// to it.  For BatchSize > 1 the reference lays its workspace out per batch but multiplies it as one K x (B*P) matrix
// (conv.go:107-137), which scrambles columns; here the column matrix is built as the GEMM expects it.
// Memory plan: the column matrix is written once and read once by the GEMM (K*B*P*8 bytes of HBM traffic, the dominant
// term); placement + bias is one pass (the reference copies C to a temp, reorders, then adds bias: three passes).
//
// Implicit GEMM (default in the exact SGEMM mode): conv_igemm_kernel never materialises the column matrix.  Its B-tile loader
// gathers col[k][n] straight from the NCHW input (a per-k table holds the (c, kh, kw) offsets, each thread owns one output
// position n for the whole K walk, so consecutive threads read consecutive input pixels), and its epilogue adds the bias and
// stores NCHW directly.  HBM traffic drops from |in| + 2*K*N*4 + 2*|out| (+ 2*|out| placement) to |in| (through L2) + |out|;
// the FMA chain per output element is unchanged, so the two routes are bit-identical (tests/test_gpu_conv.py checks both).
#include "gemm_simt.cuh"
#include <cstdlib>
#include <mutex>
#include <utility>
#include <vector>

namespace guda {

int sgemm_mode();

struct ConvGeom {
  int batch, in_c, in_h, in_w, out_c, kh, kw, stride_h, stride_w, pad_h, pad_w, dil_h, dil_w, out_h, out_w;
};

// col[row][n], row = (c*KH + kh)*KW + kw, n = b*P + h*OW + w  (conv.go:200-222); consecutive threads walk n
__global__ void __launch_bounds__(256) im2col_kernel(const float* __restrict__ in, float* __restrict__ col, ConvGeom g) {
  const long long P = (long long)g.out_h * g.out_w, N = P * g.batch;
  const long long total = (long long)g.in_c * g.kh * g.kw * N;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const long long row = e / N, n = e - row * N;
    const int kw = (int)(row % g.kw), kh = (int)((row / g.kw) % g.kh), c = (int)(row / ((long long)g.kw * g.kh));
    const int b = (int)(n / P);
    const long long p = n - (long long)b * P;
    const int h = (int)(p / g.out_w), w = (int)(p - (long long)h * g.out_w);
    const int ih = h * g.stride_h - g.pad_h + kh * g.dil_h, iw = w * g.stride_w - g.pad_w + kw * g.dil_w;
    float v = 0.0f;
    if (ih >= 0 && ih < g.in_h && iw >= 0 && iw < g.in_w) v = __ldg(in + (((long long)b * g.in_c + c) * g.in_h + ih) * g.in_w + iw);
    col[e] = v;
  }
}

// out[b][oc][p] = tmp[oc][b*P + p] (+ bias[oc])   (conv.go:141-166 in one pass); tmp == nullptr: bias added in place (B == 1)
__global__ void __launch_bounds__(256) conv_place_kernel(const float* __restrict__ tmp, const float* __restrict__ bias, float* __restrict__ out,
                                                        int batch, int out_c, long long P) {
  const long long total = (long long)batch * out_c * P;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const long long p = e % P;
    const int oc = (int)((e / P) % out_c), b = (int)(e / (P * out_c));
    float v = tmp ? tmp[(long long)oc * (P * batch) + (long long)b * P + p] : out[e];
    if (bias) v = __fadd_rn(v, bias[oc]);
    out[e] = v;
  }
}

// ---------------------------------------------------------------- implicit GEMM
// ktab[k] = {offset of (c, kh*dil_h, kw*dil_w) inside one image of the input, (kh*dil_h << 16) | kw*dil_w}
__global__ void conv_ktab_kernel(int2* __restrict__ ktab, int K, ConvGeom g) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const int kw = k % g.kw, kh = (k / g.kw) % g.kh, c = k / (g.kw * g.kh);
  const int dh = kh * g.dil_h, dw = kw * g.dil_w;
  ktab[k] = make_int2((c * g.in_h + dh) * g.in_w + dw, (dh << 16) | dw);
}

// C tile BM (output channels) x BN (output positions n = b*P + p), 256 threads as 16 x 16, micro-tile (BM/16) x (BN/16) in
// halves like gemm_simt_kernel; same chain: acc = 0; for k ascending: if (w != 0) acc = fma(w, col, acc); then + bias.
// TXN threads along n, 256 / TXN along m.  64 x 128 uses 32 x 8 (micro-tile 8 x 4, a warp = one m-row of threads): its A
// fragments are warp-wide broadcasts and its B fragments 8-byte loads, ~6-8 shared-memory wavefronts per k step where the
// 16 x 16 grid's 4 x 8 micro-tile needed ~10-12 for the same 8 SM-cycles of FFMA2 (the LSU is the co-limit).
template <int BM, int BN, bool FASTA, int MINB, int TXN>
__global__ void __launch_bounds__(256, MINB) conv_igemm_kernel(const float* __restrict__ w, const float* __restrict__ in,
                                                            const float* __restrict__ bias, float* __restrict__ out,
                                                            const int2* __restrict__ ktab, ConvGeom g, int K, long long N, int P) {
  constexpr int BK = 16, TYN = 256 / TXN, TM = BM / TYN, TN = BN / TXN, HM = TM / 2, HN = TN / 2;
  constexpr int BPT = BK * BN / 256;                 // B elements per thread per k-tile
  constexpr int KSTEP = 256 / BN;                    // k rows covered by one pass of the CTA
  static_assert(256 % BN == 0 && TN % 2 == 0 && HM >= 1, "tile shape");
  using Smem = GemmSmem<float, BM, BN, BK>;
  __shared__ __align__(16) Smem sm;

  const int tid = threadIdx.x, tx = tid % TXN, ty = tid / TXN;
  const int m0 = blockIdx.y * BM;
  const long long n0 = (long long)blockIdx.x * BN;

  // this thread's column of the B tile: one output position for the whole K walk
  const int bn = tid % BN, bk0 = tid / BN;
  const long long nn = n0 + bn;
  const bool n_ok = nn < N;
  int ih0 = 0, iw0 = 0;
  const float* img = in;
  if (n_ok) {
    const int b = (int)(nn / P), p = (int)(nn - (long long)b * P);
    const int oh = p / g.out_w, ow = p - oh * g.out_w;
    ih0 = oh * g.stride_h - g.pad_h;
    iw0 = ow * g.stride_w - g.pad_w;
    img = in + (long long)b * g.in_c * g.in_h * g.in_w + (long long)ih0 * g.in_w + iw0;
  }
  float breg[BPT];
  auto load_b = [&](int k0) {
#pragma unroll
    for (int it = 0; it < BPT; it++) {
      const int k = k0 + bk0 + it * KSTEP;
      float v = 0.0f;
      if (n_ok && k < K) {
        const int2 e = __ldg(ktab + k);
        const int ih = ih0 + (e.y >> 16), iw = iw0 + (e.y & 0xffff);
        if ((unsigned)ih < (unsigned)g.in_h && (unsigned)iw < (unsigned)g.in_w) v = __ldg(img + e.x);
      }
      breg[it] = v;
    }
  };
  auto store_b = [&](int buf) {
#pragma unroll
    for (int it = 0; it < BPT; it++) sm.b[buf][bk0 + it * KSTEP][bn] = breg[it];
  };

  TileLoader<float, BM, BK, true, FASTA> la;
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; i++)
#pragma unroll
    for (int j = 0; j < TN; j++) acc[i][j] = 0.0f;

  const int KT = (K + BK - 1) / BK;
  la.load(w, K, m0, 0, g.out_c, K, tid);
  load_b(0);
  bool z = la.template store<Smem::LDA_S>(sm.a[0], 1.0f, true, tid);
  store_b(0);
  bool zf = __syncthreads_or(z);
  for (int kt = 0; kt < KT; kt++) {
    const int cur = kt & 1;
    if (kt + 1 < KT) {
      la.load(w, K, m0, (kt + 1) * BK, g.out_c, K, tid);
      load_b((kt + 1) * BK);
    }
#pragma unroll
    for (int kk = 0; kk < BK; kk++) {
      float af[TM], bf[TN];
      load_frag<float, HM>(af, &sm.a[cur][kk][ty * HM]);
      load_frag<float, HM>(af + HM, &sm.a[cur][kk][BM / 2 + ty * HM]);
      load_frag<float, HN>(bf, &sm.b[cur][kk][tx * HN]);
      load_frag<float, HN>(bf + HN, &sm.b[cur][kk][BN / 2 + tx * HN]);
      if (!zf) {
#pragma unroll
        for (int i = 0; i < TM; i++) {
          const unsigned long long aa = dup_f32(af[i]);
#pragma unroll
          for (int j = 0; j < TN; j += 2) fma2_f32(acc[i][j], acc[i][j + 1], aa, bf[j], bf[j + 1]);
        }
      } else {
#pragma unroll
        for (int i = 0; i < TM; i++) {
          if (af[i] != 0.0f) {                                           // compute/sgemm.go:301 `if tmp != 0` (true for NaN)
#pragma unroll
            for (int j = 0; j < TN; j++) acc[i][j] = __fmaf_rn(af[i], bf[j], acc[i][j]);
          }
        }
      }
    }
    z = false;
    if (kt + 1 < KT) {
      z = la.template store<Smem::LDA_S>(sm.a[cur ^ 1], 1.0f, true, tid);
      store_b(cur ^ 1);
    }
    zf = __syncthreads_or(z);
  }

  // ---- epilogue: + bias (conv.go:160-166), NCHW store out[b][oc][p]
#pragma unroll
  for (int jh = 0; jh < 2; jh++) {
    const long long c0 = n0 + (jh == 0 ? tx * HN : BN / 2 + tx * HN);
    if (c0 >= N) continue;
    const int b0 = (int)(c0 / P), p0 = (int)(c0 - (long long)b0 * P);
#pragma unroll
    for (int i = 0; i < TM; i++) {
      const int oc = m0 + (i < HM ? ty * HM + i : BM / 2 + ty * HM + (i - HM));
      if (oc >= g.out_c) continue;
      const float bv = bias ? bias[oc] : 0.0f;
      float o[HN];
#pragma unroll
      for (int j = 0; j < HN; j++) o[j] = bias ? __fadd_rn(acc[i][jh * HN + j], bv) : acc[i][jh * HN + j];
      float* dst = out + ((long long)b0 * g.out_c + oc) * P + p0;
      if (HN % 4 == 0 && p0 + HN <= P && c0 + HN <= N && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
        for (int j = 0; j < HN; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]);
      } else if (HN == 2 && p0 + HN <= P && c0 + HN <= N && (reinterpret_cast<uintptr_t>(dst) & 7) == 0) {
        *reinterpret_cast<float2*>(dst) = make_float2(o[0], o[1]);
      } else {
        int b = b0, p = p0;
#pragma unroll
        for (int j = 0; j < HN; j++) {
          if (c0 + j < N) out[((long long)b * g.out_c + oc) * P + p] = o[j];
          if (++p == P) { p = 0; b++; }
        }
      }
    }
  }
}

// conv.go:231-300 Conv2DDirect: per output element sum = 0; for (ic, kh, kw) ascending, in-bounds taps only:
// sum = fl(sum + fl(in * w)) -- no FMA, no zero test on the weights -- then + bias.  One thread per output element walks the
// reference's loops (consecutive threads are consecutive ow, so the input reads coalesce and the weight reads broadcast).
// This entry point exists for parity with the reference's cross-check API; the fast route is Conv2D above.
__global__ void __launch_bounds__(256) conv_direct_kernel(const float* __restrict__ in, const float* __restrict__ w,
                                                         const float* __restrict__ bias, float* __restrict__ out, ConvGeom g) {
  const long long P = (long long)g.out_h * g.out_w, total = P * g.out_c * g.batch;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const long long p = e % P;
    const int oc = (int)((e / P) % g.out_c), b = (int)(e / (P * g.out_c));
    const int oh = (int)(p / g.out_w), ow = (int)(p - (long long)oh * g.out_w);
    const float* img = in + (long long)b * g.in_c * g.in_h * g.in_w;
    const float* wk = w + (long long)oc * g.in_c * g.kh * g.kw;
    float sum = 0.0f;
    for (int ic = 0; ic < g.in_c; ic++)
      for (int kh = 0; kh < g.kh; kh++) {
        const int ih = oh * g.stride_h - g.pad_h + kh * g.dil_h;
        if (ih < 0 || ih >= g.in_h) continue;
        for (int kw = 0; kw < g.kw; kw++) {
          const int iw = ow * g.stride_w - g.pad_w + kw * g.dil_w;
          if (iw < 0 || iw >= g.in_w) continue;
          sum = __fadd_rn(sum, __fmul_rn(__ldg(img + ((long long)ic * g.in_h + ih) * g.in_w + iw), __ldg(wk + (ic * g.kh + kh) * g.kw + kw)));
        }
      }
    out[e] = bias ? __fadd_rn(sum, bias[oc]) : sum;
  }
}

struct KtabKey {
  int in_c, in_h, in_w, kh, kw, dil_h, dil_w;
  bool operator==(const KtabKey& o) const {
    return in_c == o.in_c && in_h == o.in_h && in_w == o.in_w && kh == o.kh && kw == o.kw && dil_h == o.dil_h && dil_w == o.dil_w;
  }
};
static std::mutex g_ktab_mu;
static std::vector<std::pair<KtabKey, int2*>> g_ktabs;

static int conv_route_env() {          // GUDA_CONV_ROUTE=im2col forces the materialised route (A/B and parity tests)
  static const int r = [] { const char* e = getenv("GUDA_CONV_ROUTE"); return (e && !strcmp(e, "im2col")) ? 1 : 0; }();
  return r;
}
static int g_conv_route = -1;          // -1: environment, 0: implicit, 1: im2col, 2/3/4: implicit with the 64x128 / 64x64 / 128x128 tile


}  // namespace guda

using namespace guda;
extern "C" int guda_conv2d(const float* input, const float* kernel, const float* bias, float* output, const guda_conv_params_t* cp,
                           guda_stream_t stream) {
  if (!cp) return set_error(GUDA_ERR_INVALID_ARG, "Conv2D: null parameters");
  // ConvParams.Validate, conv.go:34-51 (same messages)
  if (cp->batch_size <= 0 || cp->in_channels <= 0 || cp->in_height <= 0 || cp->in_width <= 0)
    return set_error(GUDA_ERR_INVALID_ARG, "invalid input dimensions: batch, channels, height and width must be positive");
  if (cp->out_channels <= 0 || cp->kernel_height <= 0 || cp->kernel_width <= 0)
    return set_error(GUDA_ERR_INVALID_ARG, "invalid kernel dimensions: output channels and kernel size must be positive");
  if (cp->stride_h <= 0 || cp->stride_w <= 0) return set_error(GUDA_ERR_INVALID_ARG, "invalid stride: stride must be positive");
  if (cp->dilation_h <= 0 || cp->dilation_w <= 0) return set_error(GUDA_ERR_INVALID_ARG, "invalid dilation: dilation must be positive");
  if (cp->pad_h < 0 || cp->pad_w < 0) return set_error(GUDA_ERR_INVALID_ARG, "invalid padding: padding cannot be negative");
  ConvGeom g{cp->batch_size, cp->in_channels, cp->in_height, cp->in_width, cp->out_channels, cp->kernel_height, cp->kernel_width,
             cp->stride_h, cp->stride_w, cp->pad_h, cp->pad_w, cp->dilation_h, cp->dilation_w, 0, 0};
  g.out_h = (g.in_h + 2 * g.pad_h - ((g.kh - 1) * g.dil_h + 1)) / g.stride_h + 1;       // conv.go:54-63
  g.out_w = (g.in_w + 2 * g.pad_w - ((g.kw - 1) * g.dil_w + 1)) / g.stride_w + 1;
  if (g.out_h <= 0 || g.out_w <= 0) return set_error(GUDA_ERR_INVALID_ARG, "Conv2D: kernel larger than the padded input");
  int rc = bind_device();
  if (rc) return rc;
  if (!input || !kernel || !output) return set_error(GUDA_ERR_INVALID_ARG, "Conv2D: null pointer");
  const long long P = (long long)g.out_h * g.out_w, N = P * g.batch, K = (long long)g.in_c * g.kh * g.kw;
  if (N > 0x7fffffffLL || K > 0x7fffffffLL) return set_error(GUDA_ERR_INVALID_ARG, "Conv2D: problem too large for 32-bit GEMM dimensions");
  const float* b = (cp->use_bias && bias) ? bias : nullptr;
  cudaStream_t s = native(stream);
  const int route = g_conv_route >= 0 ? g_conv_route : conv_route_env();
  // the 4^3 / 8^3 / 16^3 contiguous shapes take Sgemm's unfused small-matrix kernels (compute/sgemm.go:80-96), and the 3xTF32
  // mode is a property of guda_sgemm: both stay on the materialised route
  const bool small_path = g.out_c == N && N == K && (K == 4 || K == 8 || K == 16);
  const long long in_image = (long long)g.in_c * g.in_h * g.in_w;
  // few output positions (N = B*P below 4096: at most 64 column tiles, deep K walks per CTA): the materialised route's
  // 128-bit B loads beat the scalar gather (B1 64x56x56->64: 63 vs 82 us; B32 512x7x7->512: 0.41 vs 0.53 ms); same bits either way
  const bool few_positions = route == 0 && N < 4096;
  if (route != 1 && !few_positions && !small_path && sgemm_mode() == GUDA_SGEMM_EXACT && in_image < 0x7fffffffLL && P < 0x7fffffffLL &&
      g.kh * (long long)g.dil_h < 32768 && g.kw * (long long)g.dil_w < 65536) {
    // The k table depends only on (IC, H, W, KH, KW, dilation): it is built once per geometry and kept for the library's
    // lifetime (a few KB each, at most 64 geometries), so a repeated call is ONE stream-ordered kernel launch with no
    // allocation and no synchronisation.  Beyond 64 geometries the table is a pooled temporary and the call blocks.
    PoolTemp tab;
    const int2* kt = nullptr;
    {
      std::lock_guard<std::mutex> lk(g_ktab_mu);
      const KtabKey key{g.in_c, g.in_h, g.in_w, g.kh, g.kw, g.dil_h, g.dil_w};
      for (const auto& e : g_ktabs) if (e.first == key) { kt = e.second; break; }
      if (!kt && g_ktabs.size() < 64) {
        int2* p = nullptr;
        GUDA_CUDA(cudaMalloc(&p, (size_t)K * sizeof(int2)));
        conv_ktab_kernel<<<(unsigned)((K + 127) / 128), 128, 0, s>>>(p, (int)K, g);
        if ((rc = check_launch("conv_ktab"))) { cudaFree(p); return rc; }
        GUDA_CUDA(cudaStreamSynchronize(s));      // once per geometry: later calls may come from any stream
        g_ktabs.emplace_back(key, p);
        kt = p;
      }
    }
    if (!kt) {
      if ((rc = tab.alloc((size_t)K * sizeof(int2), s))) return rc;
      conv_ktab_kernel<<<(unsigned)((K + 127) / 128), 128, 0, s>>>((int2*)tab.p, (int)K, g);
      if ((rc = check_launch("conv_ktab"))) return rc;
      kt = (const int2*)tab.p;
    }
    // CTA tile 128x128 (8x8 micro-tile, 16 x 16 threads), 64x128 (8x4, 32 x 8 threads) or 64x64 (4x4): like Sgemm's choice
    // (gemm.cu launch_simt_one), the smallest estimated makespan ceil(tiles / SMs) * BM*BN / efficiency.  A 64x256 tile (4x16
    // micro-tile, 122 registers, 16 gathers per thread per k-tile) was measured at 15 TFLOP/s against 28 for 64x128 and is not built.
    const long long sms = sm_count();
    auto makespan = [&](int bm, int bn, double eff) {
      const long long tiles = ((g.out_c + bm - 1) / bm) * ((N + bn - 1) / bn);
      return (double)((tiles + sms - 1) / sms) * bm * bn / eff;
    };
    int bm = 64, bn = 128;
    if (route == 2) { bm = 64; bn = 128; }              // A/B and test knob (guda_conv_set_route)
    else if (route == 3) { bm = 64; bn = 64; }
    else if (route == 4) { bm = 128; bn = 128; }
    else {
      const double t_big = makespan(128, 128, 1.0), t_mid = makespan(64, 128, 0.85), t_small = makespan(64, 64, 0.6);
      const long long big_tiles = ((g.out_c + 127) / 128) * ((N + 127) / 128);
      // the 8x8 tile needs two resident CTAs per SM to hide its gathers (measured: 98 tiles on 148 SMs lose to 64x64)
      if (big_tiles >= 2 * sms && t_big < t_mid && t_big <= t_small) { bm = 128; bn = 128; }
      else if (t_mid <= t_small) { bm = 64; bn = 128; }
      else { bm = 64; bn = 64; }
    }
    dim3 grid((unsigned)((N + bn - 1) / bn), (unsigned)((g.out_c + bm - 1) / bm));
    if (grid.y > 65535) return set_error(GUDA_ERR_INVALID_ARG, "Conv2D: too many output channels for one launch");
    const bool fasta = (g.out_c % bm == 0) && (K % 16 == 0) && ((uintptr_t)kernel % 16 == 0);
#define GUDA_CONV_LAUNCH(BM_, BN_, MINB_, TXN_)                                                                                     \
    do {                                                                                                                        \
      if (fasta) conv_igemm_kernel<BM_, BN_, true, MINB_, TXN_><<<grid, 256, 0, s>>>(kernel, input, b, output, kt, g, (int)K, N, (int)P);  \
      else       conv_igemm_kernel<BM_, BN_, false, MINB_, TXN_><<<grid, 256, 0, s>>>(kernel, input, b, output, kt, g, (int)K, N, (int)P); \
    } while (0)
    if (bm == 128) GUDA_CONV_LAUNCH(128, 128, 2, 16);
    else if (bn == 128) GUDA_CONV_LAUNCH(64, 128, 3, 32);
    else GUDA_CONV_LAUNCH(64, 64, 4, 16);
#undef GUDA_CONV_LAUNCH
    if ((rc = check_launch("conv_igemm"))) return rc;
    return GUDA_OK;
  }
  PoolTemp col, tmp;
  if ((rc = col.alloc((size_t)(K * N) * sizeof(float), s))) return rc;             // conv.go:88-93 workspace
  const bool direct_out = (g.batch == 1);                                                // [OC][P] already is NCHW
  if (!direct_out && (rc = tmp.alloc((size_t)g.out_c * N * sizeof(float), s))) return rc;
  long long blocks = (K * N + 255) / 256;
  if (blocks > (long long)sm_count() * 32) blocks = (long long)sm_count() * 32;
  im2col_kernel<<<(unsigned)blocks, 256, 0, s>>>(input, (float*)col.p, g);
  if ((rc = check_launch("im2col"))) return rc;
  float* gemm_out = direct_out ? output : (float*)tmp.p;
  if ((rc = guda_sgemm(0, 0, g.out_c, (int)N, (int)K, 1.0f, kernel, (int)K, (const float*)col.p, (int)N, 0.0f, gemm_out, (int)N, stream)))
    return rc;
  if (!direct_out || b) {
    long long total = (long long)g.batch * g.out_c * P;
    blocks = (total + 255) / 256;
    if (blocks > (long long)sm_count() * 32) blocks = (long long)sm_count() * 32;
    conv_place_kernel<<<(unsigned)blocks, 256, 0, s>>>(direct_out ? nullptr : (const float*)tmp.p, b, output, g.batch, g.out_c, P);
    if ((rc = check_launch("conv_place"))) return rc;
  }
  // the workspace returns to the pool here (conv.go:93 defer Free); the pool orders its next owner after this stream's work
  return GUDA_OK;
}

/* route switch for A/B measurements and the route-equality tests: 0 implicit GEMM (default), 1 materialised im2col,
 * 2 / 3 / 4 implicit GEMM with the 64x128 / 64x64 / 128x128 CTA tile forced, -1 environment */
extern "C" int guda_conv_set_route(int route) {
  if (route < -1 || route > 4) return set_error(GUDA_ERR_INVALID_ARG, "Conv2D: route must be -1 .. 4");
  g_conv_route = route;
  return GUDA_OK;
}

/* conv.go:231-300 Conv2DDirect (same validation and output geometry as Conv2D; stream-ordered) */
extern "C" int guda_conv2d_direct(const float* input, const float* kernel, const float* bias, float* output, const guda_conv_params_t* cp,
                                  guda_stream_t stream) {
  if (!cp) return set_error(GUDA_ERR_INVALID_ARG, "Conv2DDirect: null parameters");
  if (cp->batch_size <= 0 || cp->in_channels <= 0 || cp->in_height <= 0 || cp->in_width <= 0)
    return set_error(GUDA_ERR_INVALID_ARG, "invalid input dimensions: batch, channels, height and width must be positive");
  if (cp->out_channels <= 0 || cp->kernel_height <= 0 || cp->kernel_width <= 0)
    return set_error(GUDA_ERR_INVALID_ARG, "invalid kernel dimensions: output channels and kernel size must be positive");
  if (cp->stride_h <= 0 || cp->stride_w <= 0) return set_error(GUDA_ERR_INVALID_ARG, "invalid stride: stride must be positive");
  if (cp->dilation_h <= 0 || cp->dilation_w <= 0) return set_error(GUDA_ERR_INVALID_ARG, "invalid dilation: dilation must be positive");
  if (cp->pad_h < 0 || cp->pad_w < 0) return set_error(GUDA_ERR_INVALID_ARG, "invalid padding: padding cannot be negative");
  ConvGeom g{cp->batch_size, cp->in_channels, cp->in_height, cp->in_width, cp->out_channels, cp->kernel_height, cp->kernel_width,
             cp->stride_h, cp->stride_w, cp->pad_h, cp->pad_w, cp->dilation_h, cp->dilation_w, 0, 0};
  g.out_h = (g.in_h + 2 * g.pad_h - ((g.kh - 1) * g.dil_h + 1)) / g.stride_h + 1;
  g.out_w = (g.in_w + 2 * g.pad_w - ((g.kw - 1) * g.dil_w + 1)) / g.stride_w + 1;
  if (g.out_h <= 0 || g.out_w <= 0) return set_error(GUDA_ERR_INVALID_ARG, "Conv2DDirect: kernel larger than the padded input");
  int rc = bind_device();
  if (rc) return rc;
  if (!input || !kernel || !output) return set_error(GUDA_ERR_INVALID_ARG, "Conv2DDirect: null pointer");
  if ((long long)g.in_c * g.kh * g.kw > 0x7fffffffLL) return set_error(GUDA_ERR_INVALID_ARG, "Conv2DDirect: kernel too large");
  const long long total = (long long)g.batch * g.out_c * g.out_h * g.out_w;
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)sm_count() * 32) blocks = (long long)sm_count() * 32;
  conv_direct_kernel<<<(unsigned)blocks, 256, 0, native(stream)>>>(input, kernel, (cp->use_bias && bias) ? bias : nullptr, output, g);
  return check_launch("conv_direct");
}
