// conv.cu -- Conv2D (NCHW) as im2col -> exact-order SGEMM -> NCHW placement + bias.  Replaces conv.go:69-226.
//
// Same arithmetic as the reference: out[b][oc][p] = fl( chain_k fma(w[oc][k], col[k][b*P+p]) + bias[oc] ), the chain being
// Sgemm(NoTrans, NoTrans, OC, B*P, K, 1, kernel, K, col, B*P, 0, ...) of compute/sgemm.go with k = (c, kh, kw) ascending
// and the reference's zero test on the weights -- so for BatchSize == 1 (all the reference's tests) results are bit-identical
// to it.  For BatchSize > 1 the reference lays its workspace out per batch but multiplies it as one K x (B*P) matrix
// (conv.go:107-137), which scrambles columns; here the column matrix is built as the GEMM expects it.
// Memory plan: the column matrix is written once and read once by the GEMM (K*B*P*8 bytes of HBM traffic, the dominant
// term); placement + bias is one pass (the reference copies C to a temp, reorders, then adds bias: three passes).
#include "common.cuh"

namespace guda {

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

struct ConvTemp {
  void* p = nullptr;
  ~ConvTemp() { if (p) guda_free(p); }
};

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
  ConvTemp col, tmp;
  if ((rc = guda_malloc(&col.p, (size_t)(K * N) * sizeof(float)))) return rc;             // conv.go:88-93 workspace
  const bool direct_out = (g.batch == 1);                                                // [OC][P] already is NCHW
  if (!direct_out && (rc = guda_malloc(&tmp.p, (size_t)g.out_c * N * sizeof(float)))) return rc;
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
  GUDA_CUDA(cudaStreamSynchronize(s));          // the workspace returns to the pool when this function exits (conv.go:93 defer Free)
  return GUDA_OK;
}
