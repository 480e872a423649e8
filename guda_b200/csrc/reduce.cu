// reduce.cu -- whole-vector reductions on the DOT streaming skeleton: Sum, SumSquares, Max, Min, ArgMax, ArgMin, Isamax.
//
// Replaces reduce.go:13-56 (DevicePtr.Sum/Max/Min/ArgMax/ArgMin/SumSquares -> compute/asm/f32/sum_amd64.s and
// compute/asm/f32/reduce_avx2_amd64.s) and Isamax (compute/level1float32.go:89-130).
//
// Contract (what the oracle checks):
//   ArgMax / ArgMin / Isamax : bit-exact.  The reference's scalar loops update only on an ordered strict compare, so the answer is
//       "first index of the extreme among the non-NaN elements, or 0 when x[0] is NaN" -- a commutative (value, smallest index)
//       reduction once every thread starts from (x[0], 0): a NaN x[0] then never loses, exactly like VCOMISS + JBE.
//   Max / Min                : bit-exact, including NaN and signed-zero behaviour.  VMAXPS/VMINPS return the SECOND operand unless
//       the first is strictly greater (smaller), so a NaN is forgotten by the next number in its AVX lane and +-0 ties go to the
//       later element -- order-dependent, but only then.  One streaming pass computes the plain extreme plus a saw-a-NaN flag; when
//       there was no NaN and the extreme is not a zero tie (+0 and -0 both present) the plain extreme IS the reference's result.  Otherwise
//       (rare) an exact kernel replays the 8-lane fold with an ordered, associative segment summary
//       {EMPTY | NONAN v | HASNAN v} (v = fold of the elements after the segment's last NaN).
//   Sum / SumSquares         : fp32 pairwise tree (more accurate than the reference's 16 / 8 sequential lanes), deterministic for a
//       given n; compared under ReduceArchTolerance (tolerance_arch.go:99-106) like DOT.
//
// Memory plan: as DOT -- one persistent wave, 128-bit ld.global.nc, 4 independent vectors in flight per thread, shuffle tree,
// one partial per CTA, the last CTA (acq_rel ticket) folds the partials in index order and writes 16 bytes straight into mapped
// pinned host memory.  Algorithmic bytes: 4 per element.
#include "common.cuh"
#include <cmath>
#include <mutex>

namespace guda {

enum RedOp { RED_SUM = 0, RED_SUMSQ, RED_MAX, RED_MIN, RED_ARGMAX, RED_ARGMIN, RED_IAMAX };

struct __align__(16) RedResult {
  float v;              // value (sum, extreme, or the extreme's value for arg ops)
  unsigned flags;       // Max / Min: bit 0 a NaN was seen, bit 1 a +0 was seen, bit 2 a -0 was seen
  long long idx;        // index (arg ops)
};

constexpr int RED_THREADS = 256;
constexpr int RED_UNROLL = 4;
constexpr int RED_MAX_CTAS = 1024;
constexpr int RED_SLOTS = 8;

__device__ RedResult g_red_partials[RED_SLOTS][RED_MAX_CTAS];
__device__ unsigned int g_red_counter[RED_SLOTS];

__device__ __forceinline__ float4 red_ldg(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

template <int OP> struct RedState {
  float v;
  unsigned flags;
  long long idx;
};

template <int OP>
__device__ __forceinline__ void red_init(RedState<OP>& s, const float* x) {
  s.flags = 0;
  s.idx = 0;
  if constexpr (OP == RED_SUM || OP == RED_SUMSQ) s.v = 0.0f;
  else if constexpr (OP == RED_MAX) s.v = -INFINITY;
  else if constexpr (OP == RED_MIN) s.v = INFINITY;
  else if constexpr (OP == RED_IAMAX) s.v = fabsf(x[0]);
  else s.v = x[0];
}

// fold one element (arg ops and max/min; the sums use their own lane accumulators)
template <int OP>
__device__ __forceinline__ void red_elem(RedState<OP>& s, float e, long long i) {
  if constexpr (OP == RED_MAX) {
    if (e != e) s.flags |= 1u;
    else if (e > s.v) s.v = e;
    if (e == 0.0f) s.flags |= 2u << (__float_as_uint(e) >> 31);       // bit 1: saw +0, bit 2: saw -0
  } else if constexpr (OP == RED_MIN) {
    if (e != e) s.flags |= 1u;
    else if (e < s.v) s.v = e;
    if (e == 0.0f) s.flags |= 2u << (__float_as_uint(e) >> 31);
  } else if constexpr (OP == RED_ARGMAX) {
    if (e > s.v) { s.v = e; s.idx = i; }
  } else if constexpr (OP == RED_ARGMIN) {
    if (s.v > e) { s.v = e; s.idx = i; }
  } else if constexpr (OP == RED_IAMAX) {
    const float a = fabsf(e);
    if (a > s.v) { s.v = a; s.idx = i; }
  }
}

// merge two partial states; commutative for every op except the sums, whose callers fix the order
template <int OP>
__device__ __forceinline__ void red_merge(RedState<OP>& a, const RedState<OP>& b) {
  if constexpr (OP == RED_SUM || OP == RED_SUMSQ) {
    a.v = __fadd_rn(a.v, b.v);
  } else if constexpr (OP == RED_MAX) {
    a.flags |= b.flags;
    if (b.v > a.v) a.v = b.v;
  } else if constexpr (OP == RED_MIN) {
    a.flags |= b.flags;
    if (b.v < a.v) a.v = b.v;
  } else if constexpr (OP == RED_ARGMAX || OP == RED_IAMAX) {
    if (b.v > a.v || (b.v == a.v && b.idx < a.idx)) { a.v = b.v; a.idx = b.idx; }
  } else {
    if (b.v < a.v || (b.v == a.v && b.idx < a.idx)) { a.v = b.v; a.idx = b.idx; }
  }
}

template <int OP>
__device__ __forceinline__ RedState<OP> red_shfl_down(const RedState<OP>& s, int o) {
  RedState<OP> r;
  r.v = __shfl_down_sync(0xffffffffu, s.v, o);
  r.flags = (OP == RED_MAX || OP == RED_MIN) ? __shfl_down_sync(0xffffffffu, s.flags, o) : 0u;
  r.idx = (OP >= RED_ARGMAX) ? __shfl_down_sync(0xffffffffu, s.idx, o) : 0ll;
  return r;
}

template <int OP>
__device__ __forceinline__ void red_block(RedState<OP>& s, RedState<OP>* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) red_merge<OP>(s, red_shfl_down<OP>(s, o));
  if ((threadIdx.x & 31) == 0) smem[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    RedState<OP> t;
    if (threadIdx.x < RED_THREADS / 32) t = smem[threadIdx.x];
    else {                                              // neutral element
      t.flags = 0;
      t.idx = 0x7fffffffffffffffll;
      t.v = (OP == RED_SUM || OP == RED_SUMSQ) ? 0.0f : smem[0].v;
      if (OP >= RED_ARGMAX) t.idx = smem[0].idx;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) red_merge<OP>(t, red_shfl_down<OP>(t, o));
    s = t;
  }
}

template <int OP>
__global__ void __launch_bounds__(RED_THREADS, 4) reduce_kernel(size_t n, const float* __restrict__ x, long long incx, size_t head,
                                                                RedResult* __restrict__ partials, unsigned int* __restrict__ counter,
                                                                RedResult* __restrict__ out) {
  RedState<OP> s;
  red_init<OP>(s, x);
  float acc[RED_UNROLL][4];
#pragma unroll
  for (int u = 0; u < RED_UNROLL; u++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[u][j] = 0.0f;

  // Max / Min, vector path: umin == 0 <=> a +0 occurred; smin == INT_MIN <=> a -0; smax > 0x7F800000 or umax > 0xFF800000 <=> a NaN
  unsigned bits_umin = 0xFFFFFFFFu, bits_umax = 0u;
  int bits_smin = 0x7FFFFFFF, bits_smax = (int)0x80000000;
  if (incx == 1) {
    const size_t nvec = (n - head) / 4;
    const float4* xv = reinterpret_cast<const float4*>(x + head);
    const size_t stride = (size_t)gridDim.x * RED_THREADS * RED_UNROLL;
    for (size_t base = (size_t)blockIdx.x * RED_THREADS * RED_UNROLL + threadIdx.x; base < nvec; base += stride) {
      float4 r[RED_UNROLL];
#pragma unroll
      for (int u = 0; u < RED_UNROLL; u++) {
        const size_t i = base + (size_t)u * RED_THREADS;
        if (i < nvec) r[u] = red_ldg(xv + i);
      }
#pragma unroll
      for (int u = 0; u < RED_UNROLL; u++) {
        const size_t i = base + (size_t)u * RED_THREADS;
        if (i < nvec) {
          const float* e = reinterpret_cast<const float*>(&r[u]);
          if constexpr (OP == RED_SUM) {
#pragma unroll
            for (int j = 0; j < 4; j++) acc[u][j] = __fadd_rn(acc[u][j], e[j]);
          } else if constexpr (OP == RED_SUMSQ) {
#pragma unroll
            for (int j = 0; j < 4; j++) acc[u][j] = __fmaf_rn(e[j], e[j], acc[u][j]);
          } else if constexpr (OP == RED_MAX || OP == RED_MIN) {
            // branch-free: the extreme with NaN-ignoring max/min, and the three facts the host needs (a NaN, a +0, a -0 occurred)
            // as four integer extremes of the bit patterns -- independent accumulators instead of one compare/select chain, so the
            // arithmetic between two batches of loads stays short (this kernel is one load batch per iteration: latency-bound)
            const unsigned* b = reinterpret_cast<const unsigned*>(&r[u]);
            if constexpr (OP == RED_MAX) s.v = fmaxf(s.v, fmaxf(fmaxf(e[0], e[1]), fmaxf(e[2], e[3])));
            else s.v = fminf(s.v, fminf(fminf(e[0], e[1]), fminf(e[2], e[3])));
            bits_umin = min(bits_umin, min(min(b[0], b[1]), min(b[2], b[3])));
            bits_umax = max(bits_umax, max(max(b[0], b[1]), max(b[2], b[3])));
            bits_smin = min(bits_smin, min(min((int)b[0], (int)b[1]), min((int)b[2], (int)b[3])));
            bits_smax = max(bits_smax, max(max((int)b[0], (int)b[1]), max((int)b[2], (int)b[3])));
          } else {
            const long long i0 = (long long)(head + 4 * i);
#pragma unroll
            for (int j = 0; j < 4; j++) red_elem<OP>(s, e[j], i0 + j);
          }
        }
      }
    }
    if (blockIdx.x == 0) {                              // unaligned head and the < 4 element tail
      const size_t tail0 = head + nvec * 4;
      for (size_t k = threadIdx.x; k < head + (n - tail0); k += RED_THREADS) {
        const size_t i = k < head ? k : tail0 + (k - head);
        if constexpr (OP == RED_SUM) acc[0][0] = __fadd_rn(acc[0][0], x[i]);
        else if constexpr (OP == RED_SUMSQ) acc[0][0] = __fmaf_rn(x[i], x[i], acc[0][0]);
        else red_elem<OP>(s, x[i], (long long)i);
      }
    }
  } else {
    for (size_t i = (size_t)blockIdx.x * RED_THREADS + threadIdx.x; i < n; i += (size_t)gridDim.x * RED_THREADS) {
      const float e = x[(long long)i * incx];
      if constexpr (OP == RED_SUM) acc[0][0] = __fadd_rn(acc[0][0], e);
      else if constexpr (OP == RED_SUMSQ) acc[0][0] = __fmaf_rn(e, e, acc[0][0]);
      else red_elem<OP>(s, e, (long long)i);
    }
  }
  if constexpr (OP == RED_SUM || OP == RED_SUMSQ) {     // fixed-order in-thread tree
    float t = 0.0f;
#pragma unroll
    for (int u = 0; u < RED_UNROLL; u++) t = __fadd_rn(t, __fadd_rn(__fadd_rn(acc[u][0], acc[u][1]), __fadd_rn(acc[u][2], acc[u][3])));
    s.v = t;
  }

  if constexpr (OP == RED_MAX || OP == RED_MIN) {
    if (bits_smax > 0x7F800000 || bits_umax > 0xFF800000u) s.flags |= 1u;
    if (bits_umin == 0u) s.flags |= 2u;
    if (bits_smin == (int)0x80000000) s.flags |= 4u;
  }

  __shared__ RedState<OP> smem[RED_THREADS / 32];
  __shared__ bool is_last;
  red_block<OP>(s, smem);
  if (threadIdx.x == 0) {
    RedResult p;
    p.v = s.v; p.flags = s.flags; p.idx = s.idx;
    partials[blockIdx.x] = p;
    unsigned int ticket;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(ticket) : "l"(counter) : "memory");
    is_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    // deterministic final pass: thread t folds partials t, t+256, ... in index order, then the same block tree
    RedState<OP> t;
    red_init<OP>(t, x);
    if constexpr (OP == RED_ARGMAX || OP == RED_ARGMIN || OP == RED_IAMAX) t.idx = 0;
    for (unsigned int i = threadIdx.x; i < gridDim.x; i += RED_THREADS) {
      const RedResult* pp = partials + i;
      RedState<OP> b;
      b.v = __ldcg(&pp->v); b.flags = __ldcg(&pp->flags); b.idx = __ldcg(&pp->idx);
      red_merge<OP>(t, b);
    }
    __syncthreads();
    red_block<OP>(t, smem);
    if (threadIdx.x == 0) {
      RedResult r;
      r.v = t.v; r.flags = t.flags; r.idx = t.idx;
      *out = r;
      *counter = 0;
      __threadfence_system();                           // out lives in mapped pinned host memory
    }
  }
}

// ---------------------------------------------------------------------------------------------------- exact Max / Min replay
// compute/asm/f32/reduce_avx2_amd64.s:9-113 lane for lane.  op(a, b) = a > b ? a : b (a < b for Min) is what VMAXPS computes; a
// segment of one AVX lane's sequence is summarised as EMPTY, NONAN(v) or HASNAN(v) with v = fold of the elements after the
// segment's last NaN (NaN when none follow), and summaries compose associatively in sequence order.  One CTA: thread t owns a
// contiguous range of 8-element groups, ordered shuffle/shared-memory trees join them, thread 0 finishes the horizontal fold
// and the scalar tail.
enum { SEG_EMPTY = 0, SEG_NONAN = 1, SEG_HASNAN = 2 };
struct Seg { int st; float v; };

template <bool ISMAX> __device__ __forceinline__ float mm_op(float a, float b) { return ISMAX ? (a > b ? a : b) : (a < b ? a : b); }
template <bool ISMAX> __device__ __forceinline__ void seg_append(Seg& s, float e) {
  if (e != e) { s.st = SEG_HASNAN; s.v = e; }
  else if (s.st == SEG_EMPTY) { s.st = SEG_NONAN; s.v = e; }
  else s.v = mm_op<ISMAX>(s.v, e);                      // a NaN v (nothing after the last NaN yet) is replaced by e
}
template <bool ISMAX> __device__ __forceinline__ Seg seg_join(const Seg& l, const Seg& r) {
  if (r.st == SEG_EMPTY) return l;
  if (l.st == SEG_EMPTY || r.st == SEG_HASNAN) return r;
  Seg o;
  o.st = l.st;
  o.v = mm_op<ISMAX>(l.v, r.v);
  return o;
}

constexpr int MM_THREADS = 1024;

template <bool ISMAX>
__global__ void __launch_bounds__(MM_THREADS) minmax_exact_kernel(size_t n, const float* __restrict__ x, RedResult* __restrict__ out) {
  __shared__ Seg sm[MM_THREADS / 32][8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (n < 8) {                                          // scalar_loop from the start, acc = x[0]
    if (tid == 0) {
      float r = x[0];
      for (size_t i = 0; i < n; i++) r = mm_op<ISMAX>(r, x[i]);
      RedResult o; o.v = r; o.flags = 0; o.idx = 0;
      *out = o;
      __threadfence_system();
    }
    return;
  }
  const size_t groups = n >> 3;
  const size_t g0 = groups * (size_t)tid / MM_THREADS, g1 = groups * (size_t)(tid + 1) / MM_THREADS;
  Seg s[8];
#pragma unroll
  for (int l = 0; l < 8; l++) { s[l].st = SEG_EMPTY; s[l].v = 0.0f; }
  if (tid == 0) {
    const float x0 = x[0];                              // VBROADCASTSS: every lane starts from x[0]
#pragma unroll
    for (int l = 0; l < 8; l++) seg_append<ISMAX>(s[l], x0);
  }
  for (size_t g = g0; g < g1; g++) {
#pragma unroll
    for (int l = 0; l < 8; l++) seg_append<ISMAX>(s[l], x[8 * g + l]);
  }
  // ordered tree inside the warp: after step o, lane i holds segments i .. i+2o-1 (only lane 0's total is used)
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
    for (int l = 0; l < 8; l++) {
      Seg r;
      r.st = __shfl_down_sync(0xffffffffu, s[l].st, o);
      r.v = __shfl_down_sync(0xffffffffu, s[l].v, o);
      if (lane + o < 32) s[l] = seg_join<ISMAX>(s[l], r);
    }
  }
  if (lane == 0)
#pragma unroll
    for (int l = 0; l < 8; l++) sm[warp][l] = s[l];
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int l = 0; l < 8; l++) s[l] = sm[lane][l];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
      for (int l = 0; l < 8; l++) {
        Seg r;
        r.st = __shfl_down_sync(0xffffffffu, s[l].st, o);
        r.v = __shfl_down_sync(0xffffffffu, s[l].v, o);
        if (lane + o < 32) s[l] = seg_join<ISMAX>(s[l], r);
      }
    }
    if (lane == 0) {
      float q[4];
#pragma unroll
      for (int l = 0; l < 4; l++) q[l] = mm_op<ISMAX>(s[l].v, s[l + 4].v);      // VEXTRACTF128 + VMAXPS
      q[0] = mm_op<ISMAX>(q[0], q[2]);                                           // VPSRLDQ $8 + VMAXPS
      q[1] = mm_op<ISMAX>(q[1], q[3]);
      float r = mm_op<ISMAX>(q[0], q[1]);                                        // VPSRLDQ $4 + VMAXSS
      for (size_t i = groups * 8; i < n; i++) r = mm_op<ISMAX>(r, x[i]);
      RedResult o; o.v = r; o.flags = 0; o.idx = 0;
      *out = o;
      __threadfence_system();
    }
  }
}

// ---------------------------------------------------------------------------------------------------- host side
static std::mutex g_red_mu;
static unsigned g_red_slot = 0;

template <int OP>
static int launch_reduce(size_t n, const float* x, long long incx, RedResult* host_out, cudaStream_t s, const char* what) {
  Scratch* sc;
  int rc = get_scratch(&sc);
  if (rc) return rc;
  size_t head = 0;
  bool unit = incx == 1;
  if (unit) {
    if (((uintptr_t)x & 3) != 0) return set_error(GUDA_ERR_INVALID_ARG, "%s: misaligned pointer", what);
    const uintptr_t mis = (uintptr_t)x & 15;
    head = mis ? (16 - mis) / 4 : 0;
    if (head > n) head = n;
  }
  size_t work = unit ? (n / 4 + (size_t)RED_THREADS * RED_UNROLL - 1) / ((size_t)RED_THREADS * RED_UNROLL) : (n + RED_THREADS - 1) / RED_THREADS;
  unsigned grid = work > 0 ? (unsigned)(work > 65535 ? 65535 : work) : 1u;
  unsigned wave = (unsigned)sm_count() * 4u;
  if (wave > (unsigned)RED_MAX_CTAS) wave = RED_MAX_CTAS;
  if (grid > wave) grid = wave;
  RedResult* partials;
  unsigned int* counters;
  GUDA_CUDA(cudaGetSymbolAddress((void**)&partials, g_red_partials));
  GUDA_CUDA(cudaGetSymbolAddress((void**)&counters, g_red_counter));
  std::unique_lock<std::mutex> lk(g_red_mu);
  const unsigned slot = (g_red_slot++) % RED_SLOTS;
  // the second half of the library's 256-byte mapped pinned block (DOT owns the first 128 bytes)
  RedResult* dev_alias = reinterpret_cast<RedResult*>(reinterpret_cast<char*>(sc->host_scalar_dev) + 128 + slot * 16);
  volatile RedResult* host_alias = reinterpret_cast<volatile RedResult*>(reinterpret_cast<char*>(sc->host_scalar) + 128 + slot * 16);
  reduce_kernel<OP><<<grid, RED_THREADS, 0, s>>>(n, x, incx, head, partials + (size_t)slot * RED_MAX_CTAS, counters + slot, dev_alias);
  rc = check_launch(what);
  if (rc) return rc;
  GUDA_CUDA(cudaStreamSynchronize(s));
  host_out->v = host_alias->v; host_out->flags = host_alias->flags; host_out->idx = host_alias->idx;
  if (OP == RED_MAX || OP == RED_MIN) {
    // replay the reference's order only when it can matter: a NaN was seen, or the extreme is a zero AND both signs of zero occur
    // (a vector holding only +0 -- any post-ReLU tensor's Min -- or only -0 has one possible answer, the streaming pass's)
    if ((host_out->flags & 1u) || (host_out->v == 0.0f && (host_out->flags & 6u) == 6u)) {
      if (OP == RED_MAX) minmax_exact_kernel<true><<<1, MM_THREADS, 0, s>>>(n, x, dev_alias);
      else minmax_exact_kernel<false><<<1, MM_THREADS, 0, s>>>(n, x, dev_alias);
      rc = check_launch(what);
      if (rc) return rc;
      GUDA_CUDA(cudaStreamSynchronize(s));
      host_out->v = host_alias->v;
    }
  }
  return GUDA_OK;
}

template <int OP>
static int reduce_value(long long n, const float* x, float* host_result, float empty, guda_stream_t stream, const char* what) {
  if (!host_result) return set_error(GUDA_ERR_INVALID_ARG, "%s: null result pointer", what);
  *host_result = empty;
  if (n < 0) return set_error(GUDA_ERR_INVALID_ARG, "%s: negative length", what);
  if (n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!x) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", what);
  RedResult r;
  rc = launch_reduce<OP>((size_t)n, x, 1, &r, native(stream), what);
  if (rc) return rc;
  *host_result = r.v;
  return GUDA_OK;
}

template <int OP>
static int reduce_index(long long n, const float* x, long long incx, long long* host_result, guda_stream_t stream, const char* what) {
  if (!host_result) return set_error(GUDA_ERR_INVALID_ARG, "%s: null result pointer", what);
  *host_result = -1;
  if (n < 0) return set_error(GUDA_ERR_INVALID_ARG, "%s: negative length", what);
  if (n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!x) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", what);
  RedResult r;
  rc = launch_reduce<OP>((size_t)n, x, incx, &r, native(stream), what);
  if (rc) return rc;
  *host_result = r.idx;
  return GUDA_OK;
}

// ---------------------------------------------------------------------------------------------------- Dsdot / Sdsdot
// fp32 inputs, fp64 accumulation (compute/level1float32_dsdot.go, level1float32_sdsdot.go; kernels ddotunitary_amd64.s /
// ddotinc_amd64.s).  The products of two floats are exact in double, so only the summation order differs from the reference
// (pairwise here, two or four sequential lanes there): agreement to ~1e-15 relative.
__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = __dadd_rn(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}

__global__ void __launch_bounds__(RED_THREADS, 4) dsdot_kernel(size_t n, const float* __restrict__ x, long long incx, const float* __restrict__ y,
                                                               long long incy, int vec, RedResult* __restrict__ partials,
                                                               unsigned int* __restrict__ counter, RedResult* __restrict__ out) {
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  if (vec) {                                            // unit strides, both 16-byte aligned
    const size_t nvec = n / 4;
    const float4* xv = reinterpret_cast<const float4*>(x);
    const float4* yv = reinterpret_cast<const float4*>(y);
    for (size_t i = (size_t)blockIdx.x * RED_THREADS + threadIdx.x; i < nvec; i += (size_t)gridDim.x * RED_THREADS) {
      const float4 a = red_ldg(xv + i), b = red_ldg(yv + i);
      acc[0] = __fma_rn((double)a.x, (double)b.x, acc[0]);   // exact product: the FMA rounds once, like MULPD (exact) + ADDPD
      acc[1] = __fma_rn((double)a.y, (double)b.y, acc[1]);
      acc[2] = __fma_rn((double)a.z, (double)b.z, acc[2]);
      acc[3] = __fma_rn((double)a.w, (double)b.w, acc[3]);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
      const size_t i = nvec * 4 + threadIdx.x;
      acc[0] = __fma_rn((double)x[i], (double)y[i], acc[0]);
    }
  } else {
    for (size_t i = (size_t)blockIdx.x * RED_THREADS + threadIdx.x; i < n; i += (size_t)gridDim.x * RED_THREADS)
      acc[0] = __fma_rn((double)x[(long long)i * incx], (double)y[(long long)i * incy], acc[0]);
  }
  double v = __dadd_rn(__dadd_rn(acc[0], acc[1]), __dadd_rn(acc[2], acc[3]));
  v = warp_sum_f64(v);
  __shared__ double wsum[RED_THREADS / 32];
  __shared__ bool is_last;
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    double w = threadIdx.x < RED_THREADS / 32 ? wsum[threadIdx.x] : 0.0;
    w = warp_sum_f64(w);
    if (threadIdx.x == 0) {
      *reinterpret_cast<double*>(&partials[blockIdx.x]) = w;
      unsigned int ticket;
      asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(ticket) : "l"(counter) : "memory");
      is_last = (ticket == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (is_last) {
    double t = 0.0;
    for (unsigned int i = threadIdx.x; i < gridDim.x; i += RED_THREADS) t = __dadd_rn(t, __ldcg(reinterpret_cast<const double*>(&partials[i])));
    t = warp_sum_f64(t);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x < 32) {
      double w = threadIdx.x < RED_THREADS / 32 ? wsum[threadIdx.x] : 0.0;
      w = warp_sum_f64(w);
      if (threadIdx.x == 0) {
        *reinterpret_cast<double*>(out) = w;
        *counter = 0;
        __threadfence_system();
      }
    }
  }
}

static int dsdot_impl(int n, const float* x, int incx, const float* y, int incy, double* result, guda_stream_t stream, const char* what) {
  *result = 0.0;
  if (incx == 0) GUDA_PANIC("blas: zero x index increment");
  if (incy == 0) GUDA_PANIC("blas: zero y index increment");
  if (n <= 0) {
    if (n == 0) return GUDA_OK;
    GUDA_PANIC("blas: n < 0");
  }
  int rc = bind_device();
  if (rc) return rc;
  if (!x || !y) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", what);
  cudaStream_t s = native(stream);
  Scratch* sc;
  if ((rc = get_scratch(&sc))) return rc;
  const float* xs = incx < 0 ? x + (long long)(-n + 1) * incx : x;
  const float* ys = incy < 0 ? y + (long long)(-n + 1) * incy : y;
  const int vec = incx == 1 && incy == 1 && (((uintptr_t)x | (uintptr_t)y) & 15) == 0;
  size_t work = vec ? ((size_t)n / 4 + RED_THREADS - 1) / RED_THREADS : ((size_t)n + RED_THREADS - 1) / RED_THREADS;
  unsigned grid = work > 0 ? (unsigned)(work > 65535 ? 65535 : work) : 1u;
  unsigned wave = (unsigned)sm_count() * 4u;
  if (wave > (unsigned)RED_MAX_CTAS) wave = RED_MAX_CTAS;
  if (grid > wave) grid = wave;
  RedResult* partials;
  unsigned int* counters;
  GUDA_CUDA(cudaGetSymbolAddress((void**)&partials, g_red_partials));
  GUDA_CUDA(cudaGetSymbolAddress((void**)&counters, g_red_counter));
  std::unique_lock<std::mutex> lk(g_red_mu);
  const unsigned slot = (g_red_slot++) % RED_SLOTS;
  RedResult* dev_alias = reinterpret_cast<RedResult*>(reinterpret_cast<char*>(sc->host_scalar_dev) + 128 + slot * 16);
  volatile double* host_alias = reinterpret_cast<volatile double*>(reinterpret_cast<char*>(sc->host_scalar) + 128 + slot * 16);
  dsdot_kernel<<<grid, RED_THREADS, 0, s>>>((size_t)n, xs, incx, ys, incy, vec, partials + (size_t)slot * RED_MAX_CTAS, counters + slot, dev_alias);
  if ((rc = check_launch(what))) return rc;
  GUDA_CUDA(cudaStreamSynchronize(s));
  *result = *host_alias;
  return GUDA_OK;
}

// ---------------------------------------------------------------------------------------------------- Softmax / LayerNorm
// compute/asm/f32/reduce.go:152-186 fastExp, operation by operation in unfused fp32 (Go does not contract on amd64)
__device__ __forceinline__ float fast_exp_f32(float x) {
  if (x < -87.3f) return 0.0f;
  const float ln2 = 0.6931472f, inv_ln2 = 1.4426951f;
  const float sc = __fmul_rn(x, inv_ln2);
  const int k = (int)(x < 0.0f ? __fsub_rn(sc, 0.5f) : __fadd_rn(sc, 0.5f));      // Go int32(): truncation toward zero
  const float r = __fsub_rn(x, __fmul_rn((float)k, ln2));
  float p = __fadd_rn(0.041666668f, __fmul_rn(r, 0.008333334f));
  p = __fadd_rn(0.16666667f, __fmul_rn(r, p));
  p = __fadd_rn(0.5f, __fmul_rn(r, p));
  p = __fadd_rn(1.0f, __fmul_rn(r, p));
  p = __fadd_rn(1.0f, __fmul_rn(r, p));
  return __fmul_rn(p, __int_as_float((k + 127) << 23));
}

__global__ void __launch_bounds__(256) softmax_exp_kernel(size_t n, float* __restrict__ x, float mx) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    x[i] = fast_exp_f32(__fsub_rn(x[i], mx));
}

__global__ void __launch_bounds__(256) layer_norm_apply_kernel(size_t n, float* __restrict__ x, const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, float mean, float inv_std) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float nm = __fmul_rn(__fsub_rn(x[i], mean), inv_std);
    x[i] = (gamma && beta) ? __fadd_rn(__fmul_rn(nm, gamma[i]), beta[i]) : nm;
  }
}

static unsigned ew_grid(size_t n) {
  size_t g = (n + 255) / 256;
  const size_t cap = (size_t)sm_count() * 16;
  return (unsigned)(g > cap ? cap : (g ? g : 1));
}

}  // namespace guda

using namespace guda;
extern "C" {
// reduce.go:180-187 Softmax (in place over x[0..n)): Max exactly as MaxAVX2, e_i = fastExp(x_i - max) bit for bit, then the sum
// (pairwise here, one sequential chain in the reference: ReduceArchTolerance) and x_i = e_i * fl(1/sum).  Blocking.
int guda_softmax(float* x, long long n, guda_stream_t stream) {
  if (n < 0) return set_error(GUDA_ERR_INVALID_ARG, "Softmax: negative length");
  if (n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!x) return set_error(GUDA_ERR_INVALID_ARG, "Softmax: null pointer");
  cudaStream_t s = native(stream);
  RedResult r;
  if ((rc = launch_reduce<RED_MAX>((size_t)n, x, 1, &r, s, "Softmax"))) return rc;
  softmax_exp_kernel<<<ew_grid((size_t)n), 256, 0, s>>>((size_t)n, x, r.v);
  if ((rc = check_launch("Softmax"))) return rc;
  if ((rc = launch_reduce<RED_SUM>((size_t)n, x, 1, &r, s, "Softmax"))) return rc;
  const float inv = 1.0f / r.v;
  if (n > 0x7fffffffll) return set_error(GUDA_ERR_INVALID_ARG, "Softmax: n exceeds the Level-1 index range");
  if ((rc = guda_sscal((int)n, inv, x, 1, stream))) return rc;               // ScalUnitary(invSum, x)
  GUDA_CUDA(cudaStreamSynchronize(s));
  return GUDA_OK;
}

// fused.go:270-310 LayerNorm (in place; gamma and beta are applied only when BOTH are given, like the reference).  The two sums
// are pairwise (sequential in the reference: tolerance), everything after them follows the reference operation by operation.
int guda_layer_norm(float* x, const float* gamma, const float* beta, long long n, float eps, guda_stream_t stream) {
  if (n < 0) return set_error(GUDA_ERR_INVALID_ARG, "LayerNorm: negative length");
  if (n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!x) return set_error(GUDA_ERR_INVALID_ARG, "LayerNorm: null pointer");
  cudaStream_t s = native(stream);
  RedResult sum, sumsq;
  if ((rc = launch_reduce<RED_SUM>((size_t)n, x, 1, &sum, s, "LayerNorm"))) return rc;
  if ((rc = launch_reduce<RED_SUMSQ>((size_t)n, x, 1, &sumsq, s, "LayerNorm"))) return rc;
  const float fn = (float)n;
  const float mean = sum.v / fn;
  volatile float ex2 = sumsq.v / fn, m2 = mean * mean;      // volatile: keep the host compiler from fusing ex2 - mean*mean
  const float variance = ex2 - m2;
  volatile float ve = variance + eps;
  const float inv_std = 1.0f / (float)std::sqrt((double)ve);
  layer_norm_apply_kernel<<<ew_grid((size_t)n), 256, 0, s>>>((size_t)n, x, gamma, beta, mean, inv_std);
  if ((rc = check_launch("LayerNorm"))) return rc;
  GUDA_CUDA(cudaStreamSynchronize(s));
  return GUDA_OK;
}

// compute/level1float32_dsdot.go:18-54 / level1float32_sdsdot.go:18-54 (slice-length panics: host shim).  Blocking.
int guda_dsdot(int n, const float* x, int incx, const float* y, int incy, double* host_result, guda_stream_t stream) {
  if (!host_result) return set_error(GUDA_ERR_INVALID_ARG, "Dsdot: null result pointer");
  return dsdot_impl(n, x, incx, y, incy, host_result, stream, "Dsdot");
}
int guda_sdsdot(int n, float alpha, const float* x, int incx, const float* y, int incy, float* host_result, guda_stream_t stream) {
  if (!host_result) return set_error(GUDA_ERR_INVALID_ARG, "Sdsdot: null result pointer");
  double d = 0.0;
  *host_result = 0.0f;
  int rc = dsdot_impl(n, x, incx, y, incy, &d, stream, "Sdsdot");
  if (rc || n <= 0) return rc;                       // n == 0 returns 0, not alpha (level1float32_sdsdot.go:25-28)
  *host_result = alpha + (float)d;
  return GUDA_OK;
}

int guda_ssum(long long n, const float* x, float* host_result, guda_stream_t stream) { return reduce_value<RED_SUM>(n, x, host_result, 0.0f, stream, "Sum"); }
int guda_ssumsq(long long n, const float* x, float* host_result, guda_stream_t stream) { return reduce_value<RED_SUMSQ>(n, x, host_result, 0.0f, stream, "SumSquares"); }
int guda_smax(long long n, const float* x, float* host_result, guda_stream_t stream) { return reduce_value<RED_MAX>(n, x, host_result, -INFINITY, stream, "Max"); }
int guda_smin(long long n, const float* x, float* host_result, guda_stream_t stream) { return reduce_value<RED_MIN>(n, x, host_result, INFINITY, stream, "Min"); }
int guda_sargmax(long long n, const float* x, long long* host_result, guda_stream_t stream) { return reduce_index<RED_ARGMAX>(n, x, 1, host_result, stream, "ArgMax"); }
int guda_sargmin(long long n, const float* x, long long* host_result, guda_stream_t stream) { return reduce_index<RED_ARGMIN>(n, x, 1, host_result, stream, "ArgMin"); }

// compute/level1float32.go:89-130 (slice-length panic: host shim)
int guda_isamax(int n, const float* x, int incx, long long* host_result, guda_stream_t stream) {
  if (host_result) *host_result = -1;
  if (incx < 1) {
    if (incx == 0) GUDA_PANIC("blas: zero x index increment");
    return GUDA_OK;
  }
  if (n < 2) {
    if (n == 1) { if (host_result) *host_result = 0; return GUDA_OK; }
    if (n == 0) return GUDA_OK;
    GUDA_PANIC("blas: n < 0");
  }
  return reduce_index<RED_IAMAX>(n, x, incx, host_result, stream, "Isamax");
}
}

// ---------------------------------------------------------------------------------------------------- axis reductions
// reduce.go:105-178 ReduceSum / ReduceMax over a row-major m x n matrix.  Every output element is ONE sequential chain in the
// reference's index order (Sum: from float32(0), plain adds; Max: from the first element, replaced only when `val > max`), so the
// results carry the reference's bits; the parallelism is across output elements.
namespace guda {

__global__ void fill_value_kernel(float* p, size_t n, float v) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

template <bool IS_MAX> __device__ __forceinline__ float axis_op(float acc, float v) {
  if (IS_MAX) return v > acc ? v : acc;
  return __fadd_rn(acc, v);
}

// axis 0 (out[j], chain over rows): one thread per column, consecutive threads read consecutive columns; 64 rows of loads in flight
// Max only: `keep the left operand unless the right one is strictly greater` is associative, so the rows may be cut into
// gridDim.y chunks whose results are combined in chunk order by a second launch of this kernel over the [chunks][n] partials.
// Chunk 0 starts from x[0][j] like the reference (a leading NaN stays); later chunks start from -Inf, which a NaN never beats
// and which loses to nothing it should not (ties keep the earlier element, so the first of two equal zeros still wins).
template <bool IS_MAX>
__global__ void __launch_bounds__(128) reduce_axis0_kernel(const float* __restrict__ x, int m_total, int n, float* __restrict__ out,
                                                           int rows_per_chunk) {
  const long long j = (long long)blockIdx.x * 128 + threadIdx.x;
  if (j >= n) return;
  const int r0 = blockIdx.y * rows_per_chunk;
  const int m = min(rows_per_chunk, m_total - r0);
  const float* p = x + (long long)r0 * n + j;
  float acc = IS_MAX ? (blockIdx.y == 0 ? p[0] : -INFINITY) : 0.0f;
  int i = 0;                                            // Max: row 0 again is a no-op (x > x is false, also for NaN)
  for (; i + 64 <= m; i += 64) {                        // n threads in all: depth, not occupancy, has to cover the HBM latency
    float v[64];
#pragma unroll
    for (int u = 0; u < 64; u++) v[u] = __ldg(p + (long long)(i + u) * n);
#pragma unroll
    for (int u = 0; u < 64; u++) acc = axis_op<IS_MAX>(acc, v[u]);
  }
  for (; i + 16 <= m; i += 16) {
    float v[16];
#pragma unroll
    for (int u = 0; u < 16; u++) v[u] = __ldg(p + (long long)(i + u) * n);
#pragma unroll
    for (int u = 0; u < 16; u++) acc = axis_op<IS_MAX>(acc, v[u]);
  }
  for (; i < m; i++) acc = axis_op<IS_MAX>(acc, __ldg(p + (long long)i * n));
  out[(long long)blockIdx.y * n + j] = acc;
}

// axis 1 (out[i], chain over columns): a CTA owns 32 rows; all 256 threads stream 32 x 128 tiles through a double-buffered
// shared tile with coalesced (128-bit when possible) loads, and lane r of warp 0 walks row r of the tile in column order.
constexpr int AX1_ROWS = 32, AX1_COLS = 128, AX1_LD = AX1_COLS + 1;

template <bool IS_MAX>
__global__ void __launch_bounds__(256, 4) reduce_axis1_kernel(const float* __restrict__ x, int m, int n_total, float* __restrict__ out, bool vec,
                                                           int cols_per_chunk) {
  // Max only: gridDim.y column chunks (a multiple of 128 columns each), combined in chunk order by reduce_axis0_kernel<true>
  // over the [chunks][m] partials -- the same left-biased strict maximum as in the axis-0 kernel
  const int cbeg = blockIdx.y * cols_per_chunk;
  const int n = min(n_total, cbeg + cols_per_chunk);     // this CTA walks columns [cbeg, n)
  __shared__ float tile[2][AX1_ROWS][AX1_LD];
  const int tid = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * AX1_ROWS;
  const int lr = tid >> 3, lc = (tid & 7) * 4;          // this thread's row of the tile and its first column (then + 32, 64, 96)
  const long long grow = row0 + lr;
  const float* src = x + (grow < m ? grow : (long long)m - 1) * n_total;
  // Two register sets: the loads of tiles t+1 AND t+2 are in flight while warp 0 walks tile t.  With one tile ahead a CTA kept 16 KB
  // in flight (~55 KB per SM): 4.2 TB/s at the loaded DRAM latency, which is what this kernel measured (0.65 of the copy bandwidth).
  float4 reg_a[4], reg_b[4];
  auto load = [&](float4 (&reg)[4], int c0) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int c = c0 + lc + 32 * k;
      if (vec && c + 4 <= n) reg[k] = __ldg(reinterpret_cast<const float4*>(src + c));
      else {
        reg[k].x = c < n ? __ldg(src + c) : 0.0f;         reg[k].y = c + 1 < n ? __ldg(src + c + 1) : 0.0f;
        reg[k].z = c + 2 < n ? __ldg(src + c + 2) : 0.0f; reg[k].w = c + 3 < n ? __ldg(src + c + 3) : 0.0f;
      }
    }
  };
  auto store = [&](const float4 (&reg)[4], int buf) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      float* d = &tile[buf][lr][lc + 32 * k];
      d[0] = reg[k].x; d[1] = reg[k].y; d[2] = reg[k].z; d[3] = reg[k].w;
    }
  };
  const int ntiles = (n - cbeg + AX1_COLS - 1) / AX1_COLS;
  const long long myrow = row0 + tid;                    // warp 0 only
  float acc = 0.0f;
  if (IS_MAX && tid < AX1_ROWS && myrow < m) acc = blockIdx.y == 0 ? __ldg(x + myrow * n_total) : -INFINITY;
  load(reg_a, cbeg);
  store(reg_a, 0);
  if (ntiles > 1) load(reg_a, cbeg + AX1_COLS);          // tile 1
  if (ntiles > 2) load(reg_b, cbeg + 2 * AX1_COLS);      // tile 2
  __syncthreads();
  // one tile: `pend` holds tile t+1 (requested two steps ago); it is parked in the other shared buffer, then re-used for tile t+3
  auto step = [&](int t, float4 (&pend)[4]) {
    if (tid < AX1_ROWS) {
      const int cmax = min(AX1_COLS, n - cbeg - t * AX1_COLS);
      const float* r = tile[t & 1][tid];
      if (cmax == AX1_COLS) {
#pragma unroll 16
        for (int c = 0; c < AX1_COLS; c++) acc = axis_op<IS_MAX>(acc, r[c]);
      } else {
        for (int c = 0; c < cmax; c++) acc = axis_op<IS_MAX>(acc, r[c]);
      }
    }
    if (t + 1 < ntiles) store(pend, (t + 1) & 1);
    if (t + 3 < ntiles) load(pend, cbeg + (t + 3) * AX1_COLS);
    __syncthreads();
  };
  for (int t = 0; t < ntiles; t += 2) {
    step(t, reg_a);
    if (t + 1 < ntiles) step(t + 1, reg_b);
  }
  if (tid < AX1_ROWS && myrow < m) out[(long long)blockIdx.y * m + myrow] = acc;
}

static int reduce_axis_impl(const float* x, int m, int n, int axis, float* out, bool is_max, guda_stream_t stream, const char* op) {
  if (axis != 0 && axis != 1) return set_error(GUDA_ERR_INVALID_ARG, "invalid axis");                 // fused.go:14 ErrInvalidAxis
  if (m < 0 || n < 0) return set_error(GUDA_ERR_INVALID_ARG, "invalid shape");
  if ((axis == 0 ? n : m) == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!x || !out) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", op);
  if ((axis == 0 ? m : n) == 0) {
    // an empty chain: Sum leaves float32(0); Max would index x[j] out of range in the reference
    if (is_max) return set_error(GUDA_ERR_INVALID_ARG, "%s: index out of range", op);
    GUDA_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * (size_t)(axis == 0 ? n : m), native(stream)));
    return GUDA_OK;
  }
  cudaStream_t s = native(stream);
  if (axis == 0 && !is_max && m >= 64) {
    // Column sums are GemvT with alpha = 1, x = ones, beta = 0: y_j = 0, then per row y_j = fl(y_j + fl(fl(1 * 1) * a_ij)) -- the
    // same chain of plain adds in row order (1 * a is exact, NaN / Inf included), on level2.cu's warp-specialised cp.async ring
    // (6.3 instead of 1.2 TB/s at 16384^2).  The ones vector is a grow-only library-lifetime buffer.
    static std::mutex mu;
    static float* ones = nullptr;
    static size_t ones_len = 0;
    std::lock_guard<std::mutex> lk(mu);                   // held across the launch: a later grow synchronises the device before it frees
    {
      if (ones_len < (size_t)m) {
        if (ones) { GUDA_CUDA(cudaDeviceSynchronize()); GUDA_CUDA(cudaFree(ones)); ones = nullptr; ones_len = 0; }
        size_t want = 1;
        while (want < (size_t)m) want <<= 1;
        GUDA_CUDA(cudaMalloc(&ones, want * sizeof(float)));
        fill_value_kernel<<<(unsigned)((want + 255) / 256), 256, 0, s>>>(ones, want, 1.0f);
        if ((rc = check_launch(op))) return rc;
        GUDA_CUDA(cudaStreamSynchronize(s));
        ones_len = want;
      }
    }
    return guda_sgemv(1, m, n, 1.0f, x, n, ones, 1, 0.0f, out, 1, stream);
  }
  if (axis == 0) {
    const unsigned grid = (unsigned)(((long long)n + 127) / 128);
    long long chunks = 1;
    if (is_max) {                                       // enough threads to cover HBM latency: ~1024 per SM, at least 64 rows each
      chunks = ((long long)sm_count() * 1024 + n - 1) / n;
      if (chunks > m / 64) chunks = m / 64;
      if (chunks > 65535) chunks = 65535;
      if (chunks < 1) chunks = 1;
    }
    if (chunks > 1) {
      static std::mutex mu;
      static float* part = nullptr;
      static size_t part_len = 0;
      std::lock_guard<std::mutex> lk(mu);               // held across the launches: a later grow synchronises the device before it frees
      const size_t want = (size_t)chunks * (size_t)n;
      if (part_len < want) {
        if (part) { GUDA_CUDA(cudaDeviceSynchronize()); GUDA_CUDA(cudaFree(part)); part = nullptr; part_len = 0; }
        GUDA_CUDA(cudaMalloc(&part, want * sizeof(float)));
        part_len = want;
      }
      const int rpc = (int)((m + chunks - 1) / chunks);
      const unsigned ny = (unsigned)((m + rpc - 1) / rpc);
      reduce_axis0_kernel<true><<<dim3(grid, ny), 128, 0, s>>>(x, m, n, part, rpc);
      if ((rc = check_launch(op))) return rc;
      reduce_axis0_kernel<true><<<grid, 128, 0, s>>>(part, (int)ny, n, out, (int)ny);
      return check_launch(op);
    }
    if (is_max) reduce_axis0_kernel<true><<<grid, 128, 0, s>>>(x, m, n, out, m);
    else        reduce_axis0_kernel<false><<<grid, 128, 0, s>>>(x, m, n, out, m);
  } else {
    const unsigned grid = (unsigned)(((long long)m + AX1_ROWS - 1) / AX1_ROWS);
    const bool vec = ((uintptr_t)x % 16 == 0) && (n % 4 == 0);
    long long chunks = 1;
    if (is_max && grid < 2u * (unsigned)sm_count()) {   // few row blocks: cut the columns (>= 4 tiles per chunk) to fill the machine
      chunks = (4ll * sm_count() + grid - 1) / grid;
      if (chunks > n / (4 * AX1_COLS)) chunks = n / (4 * AX1_COLS);
      if (chunks > 65535) chunks = 65535;
      if (chunks < 1) chunks = 1;
    }
    if (chunks > 1) {
      static std::mutex mu;
      static float* part = nullptr;
      static size_t part_len = 0;
      std::lock_guard<std::mutex> lk(mu);
      const size_t want = (size_t)chunks * (size_t)m;
      if (part_len < want) {
        if (part) { GUDA_CUDA(cudaDeviceSynchronize()); GUDA_CUDA(cudaFree(part)); part = nullptr; part_len = 0; }
        GUDA_CUDA(cudaMalloc(&part, want * sizeof(float)));
        part_len = want;
      }
      int cpc = (int)((n + chunks - 1) / chunks);
      cpc = (cpc + AX1_COLS - 1) / AX1_COLS * AX1_COLS;
      const unsigned ny = (unsigned)((n + cpc - 1) / cpc);
      reduce_axis1_kernel<true><<<dim3(grid, ny), 256, 0, s>>>(x, m, n, part, vec, cpc);
      if ((rc = check_launch(op))) return rc;
      reduce_axis0_kernel<true><<<(unsigned)(((long long)m + 127) / 128), 128, 0, s>>>(part, (int)ny, m, out, (int)ny);
      return check_launch(op);
    }
    const int all = (n + AX1_COLS - 1) / AX1_COLS * AX1_COLS;
    if (is_max) reduce_axis1_kernel<true><<<grid, 256, 0, s>>>(x, m, n, out, vec, all);
    else        reduce_axis1_kernel<false><<<grid, 256, 0, s>>>(x, m, n, out, vec, all);
  }
  return check_launch(op);
}

}  // namespace guda

extern "C" {
int guda_reduce_sum_axis(const float* x, int m, int n, int axis, float* out, guda_stream_t stream) {
  return guda::reduce_axis_impl(x, m, n, axis, out, false, stream, "ReduceSum");
}
int guda_reduce_max_axis(const float* x, int m, int n, int axis, float* out, guda_stream_t stream) {
  return guda::reduce_axis_impl(x, m, n, axis, out, true, stream, "ReduceMax");
}
}

// ---------------------------------------------------------------------------------------------------- TopK
// reduce.go:247-291: k rounds of selection sort over (value, index) pairs -- round i scans positions i..n-1 of the CURRENT
// arrangement for the first maximum under `pairs[j].val > pairs[maxIdx].val` and swaps it into position i.  A round's scan is
// exactly the ArgMax of reduce_kernel (first index of the maximum among non-NaN elements, or the first position when it
// holds a NaN) applied to the working copy from position i on, so the indices come out bit for bit -- including the order of
// ties, which depends on the earlier swaps, and NaNs.  The permutation is tracked on the host for the k + k touched slots.
namespace guda {
__global__ void topk_swap_kernel(float* v, long long i, long long j) {
  const float t = v[i];
  v[i] = v[j];
  v[j] = t;
}
}  // namespace guda

extern "C" int guda_topk(const float* x, long long n, int k, float* host_values, long long* host_indices, guda_stream_t stream) {
  using namespace guda;
  if (n < 0) return set_error(GUDA_ERR_INVALID_ARG, "TopK: negative length");
  if (k > n) k = (int)n;                                                                   // reduce.go:248-250
  if (k <= 0) return set_error(GUDA_ERR_INVALID_ARG, "invalid parameter");                  // ErrInvalidParameter, reduce.go:251-253
  if (!host_values || !host_indices) return set_error(GUDA_ERR_INVALID_ARG, "TopK: null result pointer");
  int rc = bind_device();
  if (rc) return rc;
  if (!x) return set_error(GUDA_ERR_INVALID_ARG, "TopK: null pointer");
  cudaStream_t s = native(stream);
  PoolTemp work;
  if ((rc = work.alloc((size_t)n * sizeof(float), s))) return rc;
  float* v = static_cast<float*>(work.p);
  GUDA_CUDA(cudaMemcpyAsync(v, x, (size_t)n * sizeof(float), cudaMemcpyDeviceToDevice, s));
  std::vector<std::pair<long long, long long>> moved;                                      // position -> original index, where it differs
  auto index_at = [&](long long pos) {
    for (auto it = moved.rbegin(); it != moved.rend(); ++it) if (it->first == pos) return it->second;
    return pos;
  };
  for (int i = 0; i < k; i++) {
    RedResult r;
    if ((rc = launch_reduce<RED_ARGMAX>((size_t)(n - i), v + i, 1, &r, s, "TopK"))) return rc;  // blocks; r.idx is relative to v + i
    const long long pos = i + r.idx;
    const long long at_i = index_at(i), at_pos = index_at(pos);
    if (pos != i) {
      topk_swap_kernel<<<1, 1, 0, s>>>(v, i, pos);
      if ((rc = check_launch("TopK"))) return rc;
      moved.emplace_back(i, at_pos);
      moved.emplace_back(pos, at_i);
    }
    host_indices[i] = at_pos;
  }
  GUDA_CUDA(cudaMemcpyAsync(host_values, v, (size_t)k * sizeof(float), cudaMemcpyDeviceToHost, s));
  GUDA_CUDA(cudaStreamSynchronize(s));
  return GUDA_OK;
}
