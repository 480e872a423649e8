// elementwise.cu -- the reference's elementwise device ops as HBM-streaming kernels.
// Replaces math.go:133-169 (Add, Multiply), :230-272 (ReLU, Sigmoid -> activations.go:9-27 SigmoidFloat32) and
// fused.go:18-216 FusedKernel.Execute (one read, the op list applied to the value in registers, one write), which
// LinearReLU (fused.go:252) and GELU (:261, Custom(geluFloat32)) are built from.  The reference runs these as host closures
// over a goroutine grid (kernel.go Launch); here each is one pass of 128-bit loads/stores, algorithmic bytes 12 B/elem
// (Add, Multiply), 8 B/elem (in-place ops) + 4 B/elem per full-size operand of a fused op list.
// Arithmetic: every Go statement on a float32 rounds once and GOAMD64=v1 does not contract `val += alpha*other`, so each
// step is an explicit __fmul_rn / __fadd_rn / __fdiv_rn; ExpFloat32 and geluFloat32 are the restatements of gemm_simt.cuh.
#include "gemm_simt.cuh"
#ifndef GUDA_ACT_FAST
#define GUDA_ACT_FAST true      // false: the kernels below use the plain div.rn forms (A/B timing arm, tools/time_activations.py)
#endif

namespace guda {

template <bool FAST = true>
__device__ __forceinline__ float sigmoid_f32_exact(float x) {      // activations.go:9-27
  if (x < -10.0f) return 0.0f;
  if (x > 10.0f) return 1.0f;
  if (x >= 0.0f) {
    const float e = gelu_exp_f32(-x);
    return FAST ? div_normal_rn(1.0f, __fadd_rn(1.0f, e)) : __fdiv_rn(1.0f, __fadd_rn(1.0f, e));
  }
  const float e = gelu_exp_f32(x);                                 // also the NaN route (every comparison false): generic division
  if (!FAST || x != x) return __fdiv_rn(e, __fadd_rn(1.0f, e));
  return div_normal_rn(e, __fadd_rn(1.0f, e));
}

// All 2^32 inputs through both forms of each activation (see the comment above div_const_rn in gemm_simt.cuh)
__global__ void __launch_bounds__(256) activation_selftest_kernel(unsigned long long* counts, unsigned int* first_bad) {
  unsigned long long bad_g = 0, bad_s = 0;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < (1ull << 32); i += stride) {
    const float x = __uint_as_float((unsigned)i);
    const unsigned g1 = __float_as_uint(gelu_f32_exact<true>(x)), g0 = __float_as_uint(gelu_f32_exact<false>(x));
    const unsigned s1 = __float_as_uint(sigmoid_f32_exact<true>(x)), s0 = __float_as_uint(sigmoid_f32_exact<false>(x));
    if (g1 != g0) { bad_g++; atomicMin(first_bad, (unsigned)i); }
    if (s1 != s0) { bad_s++; atomicMin(first_bad + 1, (unsigned)i); }
  }
  if (bad_g) atomicAdd(counts, bad_g);
  if (bad_s) atomicAdd(counts + 1, bad_s);
}

enum { EW_ADD = 0, EW_MUL = 1, EW_RELU = 2, EW_SIGMOID = 3 };

// c = a (+|*) b, or x = f(x) in place (b == nullptr): 4 elements per thread per step, grid-stride
template <int OP>
__global__ void __launch_bounds__(256) ew_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ c,
                                                 size_t n, bool vec) {
  const size_t nv = vec ? n / 4 : 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  auto f = [](float x, float y) -> float {
    if (OP == EW_ADD) return __fadd_rn(x, y);
    if (OP == EW_MUL) return __fmul_rn(x, y);
    if (OP == EW_RELU) return x < 0.0f ? 0.0f : x;                 // math.go:240: NaN and -0 stay
    return sigmoid_f32_exact<GUDA_ACT_FAST>(x);
  };
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
    const float4 x = reinterpret_cast<const float4*>(a)[i];
    float4 y = make_float4(0.f, 0.f, 0.f, 0.f);
    if (OP == EW_ADD || OP == EW_MUL) y = reinterpret_cast<const float4*>(b)[i];
    reinterpret_cast<float4*>(c)[i] = make_float4(f(x.x, y.x), f(x.y, y.y), f(x.z, y.z), f(x.w, y.w));
  }
  for (size_t i = nv * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    c[i] = f(a[i], (OP == EW_ADD || OP == EW_MUL) ? b[i] : 0.0f);
}

// ---- FusedKernel.Execute (fused.go:123-216)
struct FusedArgs {
  int n_ops, n_others;
  guda_fused_op_t ops[GUDA_FUSED_MAX_OPS];
  const float* others[GUDA_FUSED_MAX_OTHERS];
};

// The op list is decoded once per group of NV values (all elements run the same list, so the switch is warp-uniform); binary
// operands are fetched with the same 128-bit loads as x when they are aligned and not broadcast.
template <int NV>
__device__ __forceinline__ void fused_apply(const FusedArgs& fa, float (&val)[NV], size_t idx0, bool vec_others) {
  int oi = 0;
  for (int o = 0; o < fa.n_ops; o++) {
    const guda_fused_op_t op = fa.ops[o];
    switch (op.type) {
      case GUDA_FUSED_ADD:
      case GUDA_FUSED_MUL:
        if (oi < fa.n_others) {                                    // fused.go:148,170: skipped when no operand is left
          float ov[NV];
          const float* other = fa.others[oi];
          if (op.broadcast_size > 0) {
#pragma unroll
            for (int j = 0; j < NV; j++) ov[j] = other[((idx0 + j) / (size_t)op.broadcast_size) % (size_t)op.broadcast_dim];
          } else if (NV % 4 == 0 && vec_others) {
#pragma unroll
            for (int j = 0; j < NV; j += 4) *reinterpret_cast<float4*>(&ov[j]) = *reinterpret_cast<const float4*>(other + idx0 + j);
          } else {
#pragma unroll
            for (int j = 0; j < NV; j++) ov[j] = other[idx0 + j];
          }
          if (op.type == GUDA_FUSED_ADD) {
#pragma unroll
            for (int j = 0; j < NV; j++) val[j] = __fadd_rn(val[j], __fmul_rn(op.alpha, ov[j]));
          } else {
#pragma unroll
            for (int j = 0; j < NV; j++) val[j] = __fmul_rn(val[j], ov[j]);
          }
          oi++;
        }
        break;
      case GUDA_FUSED_ADD_SCALAR:
#pragma unroll
        for (int j = 0; j < NV; j++) val[j] = __fadd_rn(val[j], op.alpha);
        break;
      case GUDA_FUSED_MUL_SCALAR:
#pragma unroll
        for (int j = 0; j < NV; j++) val[j] = __fmul_rn(val[j], op.alpha);
        break;
      case GUDA_FUSED_RELU:
#pragma unroll
        for (int j = 0; j < NV; j++) if (val[j] < 0.0f) val[j] = 0.0f;
        break;
      case GUDA_FUSED_SIGMOID:
#pragma unroll
        for (int j = 0; j < NV; j++) val[j] = sigmoid_f32_exact<GUDA_ACT_FAST>(val[j]);
        break;
      case GUDA_FUSED_GELU:
#pragma unroll
        for (int j = 0; j < NV; j++) val[j] = gelu_f32_exact<GUDA_ACT_FAST>(val[j]);
        break;
      default: break;
    }
  }
}

__global__ void __launch_bounds__(256) fused_execute_kernel(float* __restrict__ x, size_t n, bool vec, bool vec_others,
                                                            const __grid_constant__ FusedArgs fa) {
  const size_t n8 = vec ? n / 8 : 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float v[8];
    *reinterpret_cast<float4*>(&v[0]) = reinterpret_cast<const float4*>(x)[2 * i];
    *reinterpret_cast<float4*>(&v[4]) = reinterpret_cast<const float4*>(x)[2 * i + 1];
    fused_apply<8>(fa, v, 8 * i, vec_others);
    reinterpret_cast<float4*>(x)[2 * i] = *reinterpret_cast<float4*>(&v[0]);
    reinterpret_cast<float4*>(x)[2 * i + 1] = *reinterpret_cast<float4*>(&v[4]);
  }
  for (size_t i = n8 * 8 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float v[1] = {x[i]};
    fused_apply<1>(fa, v, i, false);
    x[i] = v[0];
  }
}

static unsigned ew_grid(size_t n) {
  size_t blocks = (n / 4 + 255) / 256;
  const size_t cap = (size_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  return (unsigned)(blocks ? blocks : 1);
}

template <int OP>
static int ew_launch(const float* a, const float* b, float* c, long long n, guda_stream_t stream, const char* op) {
  if (n < 0) return set_error(GUDA_ERR_INVALID_ARG, "%s: n < 0", op);
  if (n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!a || !c || ((OP == EW_ADD || OP == EW_MUL) && !b)) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", op);
  const bool vec = (((uintptr_t)a | (uintptr_t)c | (uintptr_t)b) & 15) == 0;
  ew_kernel<OP><<<ew_grid((size_t)n), 256, 0, native(stream)>>>(a, b, c, (size_t)n, vec);
  return check_launch(op);
}

}  // namespace guda

using namespace guda;
extern "C" {
int guda_add(const float* a, const float* b, float* c, long long n, guda_stream_t stream) { return ew_launch<EW_ADD>(a, b, c, n, stream, "Add"); }
int guda_multiply(const float* a, const float* b, float* c, long long n, guda_stream_t stream) {
  return ew_launch<EW_MUL>(a, b, c, n, stream, "Multiply");
}
int guda_relu(float* x, long long n, guda_stream_t stream) { return ew_launch<EW_RELU>(x, nullptr, x, n, stream, "ReLU"); }
int guda_sigmoid(float* x, long long n, guda_stream_t stream) { return ew_launch<EW_SIGMOID>(x, nullptr, x, n, stream, "Sigmoid"); }
// counts[0] / counts[1]: inputs (of all 2^32 bit patterns) on which the fast and the __fdiv_rn forms of geluFloat32 / SigmoidFloat32
// disagree in any bit; first_bad[0] / [1]: the smallest such input.  Blocking.  The product kernels use the fast forms.
int guda_selftest_activations(unsigned long long* counts, unsigned int* first_bad) {
  if (!counts || !first_bad) return guda::set_error(GUDA_ERR_INVALID_ARG, "selftest_activations: null pointer");
  int rc = guda::bind_device();
  if (rc) return rc;
  cudaStream_t s = guda::native(nullptr);
  guda::PoolTemp buf;
  if ((rc = buf.alloc(64, s))) return rc;
  unsigned long long* d_counts = static_cast<unsigned long long*>(buf.p);
  unsigned int* d_first = reinterpret_cast<unsigned int*>(d_counts + 2);
  GUDA_CUDA(cudaMemsetAsync(d_counts, 0, 16, s));
  GUDA_CUDA(cudaMemsetAsync(d_first, 0xFF, 8, s));
  guda::activation_selftest_kernel<<<guda::sm_count() * 8, 256, 0, s>>>(d_counts, d_first);
  if ((rc = guda::check_launch("selftest_activations"))) return rc;
  GUDA_CUDA(cudaMemcpyAsync(counts, d_counts, 16, cudaMemcpyDeviceToHost, s));
  GUDA_CUDA(cudaMemcpyAsync(first_bad, d_first, 8, cudaMemcpyDeviceToHost, s));
  GUDA_CUDA(cudaStreamSynchronize(s));
  return GUDA_OK;
}

int guda_fused_execute(float* x, const float* const* others, int n_others, long long n, const guda_fused_op_t* ops, int n_ops,
                       guda_stream_t stream) {
  if (n < 0) return set_error(GUDA_ERR_INVALID_ARG, "FusedOp: n < 0");
  if (n_ops < 0 || n_ops > GUDA_FUSED_MAX_OPS) return set_error(GUDA_ERR_INVALID_ARG, "FusedOp: at most %d ops", GUDA_FUSED_MAX_OPS);
  if (n_others < 0 || n_others > GUDA_FUSED_MAX_OTHERS)
    return set_error(GUDA_ERR_INVALID_ARG, "FusedOp: at most %d operand tensors", GUDA_FUSED_MAX_OTHERS);
  if (n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!x || (n_ops && !ops) || (n_others && !others)) return set_error(GUDA_ERR_INVALID_ARG, "FusedOp: null pointer");
  FusedArgs fa{};
  fa.n_ops = n_ops; fa.n_others = n_others;
  for (int i = 0; i < n_ops; i++) {
    fa.ops[i] = ops[i];
    if (ops[i].type < GUDA_FUSED_ADD || ops[i].type > GUDA_FUSED_GELU) return set_error(GUDA_ERR_INVALID_ARG, "FusedOp: invalid parameter");
    if (ops[i].broadcast_size > 0 && ops[i].broadcast_dim <= 0) return set_error(GUDA_ERR_INVALID_ARG, "FusedOp: invalid shape");
  }
  for (int i = 0; i < n_others; i++) {
    if (!others[i]) return set_error(GUDA_ERR_INVALID_ARG, "FusedOp: null operand tensor");
    fa.others[i] = others[i];
  }
  const bool vec = ((uintptr_t)x & 15) == 0;
  bool vec_others = true;
  for (int i = 0; i < n_others; i++) vec_others = vec_others && ((uintptr_t)others[i] & 15) == 0;
  size_t blocks = ((size_t)n / 8 + 255) / 256;
  const size_t cap = (size_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  fused_execute_kernel<<<(unsigned)(blocks ? blocks : 1), 256, 0, native(stream)>>>(x, (size_t)n, vec, vec_others, fa);
  return check_launch("FusedOp");
}
}
