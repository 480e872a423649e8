// level1.cu -- AXPY / SCAL / DOT for fp32 and fp64: HBM-streaming kernels.
//
// Replaces compute/level1float32.go:225-259 (Saxpy), :619-653 (Sscal), level1float32_sdot.go:18-54 (Sdot)
// and the fp64 twins (level1float64.go:211,595; level1float64_ddot.go:14) together with the
// assembly under compute/asm/f32 and compute/asm/f64.
//
// Arithmetic contract (what the oracle checks):
//   saxpy inc==1 : y = fma(alpha, x, y) on every element (axpyunitary_fma_amd64.s is all-FMA)  -> bit-exact
//   saxpy inc!=1 : y = (alpha*x) + y unfused (axpyinc_amd64.s MULSS/ADDSS)                     -> bit-exact
//   daxpy inc==1 : fused except the alignment peel before y is 32-byte aligned and the last
//                  n%4 elements, which are unfused (f64/axpyunitary_fma_amd64.s)               -> bit-exact
//   scal         : x *= alpha; alpha==0 -> exact zero fill; incx<0 -> no-op                     -> bit-exact
//   dot          : fp32 (fp64) FMA accumulation, pairwise tree order (more accurate than the
//                  reference's 32 sequential lanes); compared under ReduceArchTolerance.
//
// Memory plan: 128-bit ld.global.nc / st.global per thread, 4 independent vectors in flight per
// thread, one CTA per 4096-element chunk (AXPY/SCAL); DOT is one persistent wave of CTAs
// (4 per SM) with a deterministic last-block final reduction that writes the scalar straight
// into mapped pinned host memory, so the host needs no extra copy.
#include "common.cuh"
#include <map>
#include <mutex>

namespace guda {

template <typename T> struct Vec;
template <> struct Vec<float> { using type = float4; static constexpr int N = 4; };
template <> struct Vec<double> { using type = double2; static constexpr int N = 2; };

__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ double2 ldg_stream(const double2* p) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}
// y is read then overwritten by the same thread: plain (coherent) vector load, no L1 allocation
__device__ __forceinline__ float4 ld_rw(const float4* p) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ double2 ld_rw(const double2* p) {
  double2 v;
  asm volatile("ld.global.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream(float4* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_stream(double2* p, double2 v) {
  asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

__device__ __forceinline__ float fma_t(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double fma_t(double a, double b, double c) { return __fma_rn(a, b, c); }
__device__ __forceinline__ float mul_t(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_t(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float add_t(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_t(double a, double b) { return __dadd_rn(a, b); }

constexpr int L1_THREADS = 256;
constexpr int L1_UNROLL = 4;

// ---------------------------------------------------------------- AXPY, unit stride
// fused_lo/fused_hi: elements in [fused_lo, fused_hi) use FMA, the others mul+add (fp64 rule; fp32: all fused)
template <typename T>
__global__ void __launch_bounds__(L1_THREADS) axpy_unit_kernel(size_t n, T alpha, const T* __restrict__ x, T* __restrict__ y,
                                                               size_t head, size_t fused_lo, size_t fused_hi) {
  using V = typename Vec<T>::type;
  constexpr int VN = Vec<T>::N;
  // [0, head): scalar prologue up to 16-byte alignment of BOTH pointers (host guarantees same misalignment)
  const size_t nvec = (n - head) / VN;
  const V* xv = reinterpret_cast<const V*>(x + head);
  V* yv = reinterpret_cast<V*>(y + head);
  const size_t chunk = (size_t)L1_THREADS * L1_UNROLL;
  size_t base = (size_t)blockIdx.x * chunk + threadIdx.x;
  V xr[L1_UNROLL], yr[L1_UNROLL];
#pragma unroll
  for (int u = 0; u < L1_UNROLL; u++) {
    size_t i = base + (size_t)u * L1_THREADS;
    if (i < nvec) { xr[u] = ldg_stream(xv + i); yr[u] = ld_rw(yv + i); }
  }
#pragma unroll
  for (int u = 0; u < L1_UNROLL; u++) {
    size_t i = base + (size_t)u * L1_THREADS;
    if (i < nvec) {
      T* xe = reinterpret_cast<T*>(&xr[u]);
      T* ye = reinterpret_cast<T*>(&yr[u]);
      size_t e0 = head + i * VN;
#pragma unroll
      for (int j = 0; j < VN; j++) {
        bool fused = (e0 + j >= fused_lo) && (e0 + j < fused_hi);
        ye[j] = fused ? fma_t(alpha, xe[j], ye[j]) : add_t(mul_t(xe[j], alpha), ye[j]);
      }
      st_stream(yv + i, yr[u]);
    }
  }
  // scalar head and tail handled by the first CTA
  if (blockIdx.x == 0) {
    size_t tail0 = head + nvec * VN;
    for (size_t e = threadIdx.x; e < head + (n - tail0); e += L1_THREADS) {
      size_t idx = e < head ? e : tail0 + (e - head);
      bool fused = (idx >= fused_lo) && (idx < fused_hi);
      T xe = x[idx], ye = y[idx];
      y[idx] = fused ? fma_t(alpha, xe, ye) : add_t(mul_t(xe, alpha), ye);
    }
  }
}

// misaligned (x and y disagree modulo 16 bytes) or strided: scalar, one element per thread per step
template <typename T, bool kFusedAll>
__global__ void __launch_bounds__(L1_THREADS) axpy_strided_kernel(size_t n, T alpha, const T* __restrict__ x, long long incx,
                                                                  T* __restrict__ y, long long incy, size_t fused_lo, size_t fused_hi) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    T xe = x[(long long)i * incx];
    T ye = y[(long long)i * incy];
    bool fused = kFusedAll || ((i >= fused_lo) && (i < fused_hi));
    y[(long long)i * incy] = fused ? fma_t(alpha, xe, ye) : add_t(mul_t(xe, alpha), ye);
  }
}

// ---------------------------------------------------------------- SCAL
template <typename T>
__global__ void __launch_bounds__(L1_THREADS) scal_unit_kernel(size_t n, T alpha, T* __restrict__ x, size_t head, int zero) {
  using V = typename Vec<T>::type;
  constexpr int VN = Vec<T>::N;
  const size_t nvec = (n - head) / VN;
  V* xv = reinterpret_cast<V*>(x + head);
  const size_t chunk = (size_t)L1_THREADS * L1_UNROLL;
  size_t base = (size_t)blockIdx.x * chunk + threadIdx.x;
  V xr[L1_UNROLL];
  if (!zero) {
#pragma unroll
    for (int u = 0; u < L1_UNROLL; u++) {
      size_t i = base + (size_t)u * L1_THREADS;
      if (i < nvec) xr[u] = ld_rw(xv + i);
    }
  }
#pragma unroll
  for (int u = 0; u < L1_UNROLL; u++) {
    size_t i = base + (size_t)u * L1_THREADS;
    if (i < nvec) {
      T* xe = reinterpret_cast<T*>(&xr[u]);
#pragma unroll
      for (int j = 0; j < VN; j++) xe[j] = zero ? T(0) : mul_t(xe[j], alpha);
      st_stream(xv + i, xr[u]);
    }
  }
  if (blockIdx.x == 0) {
    size_t tail0 = head + nvec * VN;
    for (size_t e = threadIdx.x; e < head + (n - tail0); e += L1_THREADS) {
      size_t idx = e < head ? e : tail0 + (e - head);
      x[idx] = zero ? T(0) : mul_t(x[idx], alpha);
    }
  }
}
template <typename T>
__global__ void __launch_bounds__(L1_THREADS) scal_strided_kernel(size_t n, T alpha, T* __restrict__ x, long long incx, int zero) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    x[(long long)i * incx] = zero ? T(0) : mul_t(x[(long long)i * incx], alpha);
}

// ---------------------------------------------------------------- DOT
template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = add_t(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}

// One wave of CTAs; each thread keeps L1_UNROLL*VN independent FMA accumulators over a grid-stride
// walk of 128-bit vectors; warp shuffle tree -> smem -> one partial per CTA; the last CTA to finish
// (atomic ticket) sums the partials in index order, so the result is deterministic for a given n.
template <typename T>
__global__ void __launch_bounds__(L1_THREADS, 4) dot_kernel(size_t n, const T* __restrict__ x, long long incx, const T* __restrict__ y,
                                                         long long incy, size_t head, T* __restrict__ partials,
                                                         unsigned int* __restrict__ counter, T* __restrict__ out0, T* __restrict__ out1) {
  using V = typename Vec<T>::type;
  constexpr int VN = Vec<T>::N;
  T acc[L1_UNROLL][VN];
#pragma unroll
  for (int u = 0; u < L1_UNROLL; u++)
#pragma unroll
    for (int j = 0; j < VN; j++) acc[u][j] = T(0);
  pdl_wait();              // x / y may be the previous kernel's output; everything above ran under that kernel's tail

  const bool unit = (incx == 1 && incy == 1);
  if (unit) {
    const size_t nvec = (n - head) / VN;
    const V* xv = reinterpret_cast<const V*>(x + head);
    const V* yv = reinterpret_cast<const V*>(y + head);
    const size_t stride = (size_t)gridDim.x * L1_THREADS * L1_UNROLL;
    for (size_t base = (size_t)blockIdx.x * L1_THREADS * L1_UNROLL + threadIdx.x; base < nvec; base += stride) {
      V xr[L1_UNROLL], yr[L1_UNROLL];
#pragma unroll
      for (int u = 0; u < L1_UNROLL; u++) {
        size_t i = base + (size_t)u * L1_THREADS;
        if (i < nvec) { xr[u] = ldg_stream(xv + i); yr[u] = ldg_stream(yv + i); }
      }
#pragma unroll
      for (int u = 0; u < L1_UNROLL; u++) {
        size_t i = base + (size_t)u * L1_THREADS;
        if (i < nvec) {
          const T* xe = reinterpret_cast<const T*>(&xr[u]);
          const T* ye = reinterpret_cast<const T*>(&yr[u]);
#pragma unroll
          for (int j = 0; j < VN; j++) acc[u][j] = fma_t(xe[j], ye[j], acc[u][j]);
        }
      }
    }
    if (blockIdx.x == 0) {
      size_t tail0 = head + nvec * VN;
      for (size_t e = threadIdx.x; e < head + (n - tail0); e += L1_THREADS) {
        size_t idx = e < head ? e : tail0 + (e - head);
        acc[0][0] = fma_t(x[idx], y[idx], acc[0][0]);
      }
    }
  } else {
    for (size_t i = (size_t)blockIdx.x * L1_THREADS + threadIdx.x; i < n; i += (size_t)gridDim.x * L1_THREADS)
      acc[0][0] = fma_t(x[(long long)i * incx], y[(long long)i * incy], acc[0][0]);
  }
  pdl_trigger();           // the streaming part is over: the next kernel's CTAs may be scheduled under this kernel's fold
  // fixed-order in-thread tree
  T s = T(0);
#pragma unroll
  for (int u = 0; u < L1_UNROLL; u++) {
    T t = acc[u][0];
#pragma unroll
    for (int j = 1; j < VN; j++) t = add_t(t, acc[u][j]);
    s = add_t(s, t);
  }
  s = warp_sum(s);
  __shared__ T wsum[L1_THREADS / 32];
  __shared__ bool is_last;
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    T v = threadIdx.x < L1_THREADS / 32 ? wsum[threadIdx.x] : T(0);
    v = warp_sum(v);
    if (threadIdx.x == 0) {
      partials[blockIdx.x] = v;
      // one acq_rel ticket instead of membar + atomic: the release half publishes this CTA's partial, the acquire half
      // (in the CTA that draws the last ticket) makes every earlier partial visible -- one L2 round trip on the tail
      unsigned int ticket;
      asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(ticket) : "l"(counter) : "memory");
      is_last = (ticket == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (is_last) {
    // deterministic final pass: thread t sums partials t, t+256, ... then the same block tree
    // (up to four partials per thread are requested before the first add: one L2 round trip on the tail instead of one per partial)
    T v = T(0);
    for (unsigned int base = threadIdx.x; base < gridDim.x; base += 4 * L1_THREADS) {
      T p[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const unsigned int i = base + j * L1_THREADS;
        p[j] = i < gridDim.x ? __ldcg(partials + i) : T(0);
      }
#pragma unroll
      for (int j = 0; j < 4; j++) v = add_t(v, p[j]);
    }
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
      T w = threadIdx.x < L1_THREADS / 32 ? wsum[threadIdx.x] : T(0);
      w = warp_sum(w);
      if (threadIdx.x == 0) {
        if (out0) *out0 = w;
        if (out1) *out1 = w;
        *counter = 0;       // re-arm for the next launch on this slot
        if (out0) __threadfence_system();     // out0 lives in mapped pinned host memory
      }
    }
  }
}

// ---------------------------------------------------------------- host side
// Grid of the persistent DOT wave: at most `wave` CTAs, and as few idle chunk slots in the last round as possible -- with
// `rounds` = ceil(work / wave) every CTA walks `rounds` chunks except for a remainder smaller than `rounds` (2^24 floats: 4096
// chunks on a 592-CTA wave = 7 rounds; 586 CTAs x 7 = 4102 slots instead of 592 x 7 = 4144).
static inline size_t dot_grid(size_t work, size_t wave) {
  if (work <= wave) return work ? work : 1;
  const size_t rounds = (work + wave - 1) / wave;
  return (work + rounds - 1) / rounds;
}

template <typename T> static size_t head_elems(const void* p) {
  uintptr_t mis = (uintptr_t)p & 15;
  return mis ? (16 - mis) / sizeof(T) : 0;
}

static std::mutex g_dot_mu;
// One scratch slot (partials + ticket counter + pinned host scalar) PER STREAM: DOTs on one stream are ordered by the stream, DOTs
// on different streams (or host threads) never share partials or a ticket -- the round-1 scheme rotated 8 global slots and released
// its lock before the launch on the device-result path, so more than 8 DOTs in flight could fold each other's partials.
struct DotSlot { void* partials; unsigned int* counter; void* host_scalar; void* host_scalar_dev; };
static std::map<cudaStream_t, DotSlot> g_dot_slots;
static int dot_slot_for(cudaStream_t s, DotSlot* out) {
  std::lock_guard<std::mutex> lk(g_dot_mu);
  auto it = g_dot_slots.find(s);
  if (it == g_dot_slots.end()) {
    DotSlot d{};
    GUDA_CUDA(cudaMalloc(&d.partials, 8192));
    GUDA_CUDA(cudaMalloc((void**)&d.counter, 256));
    GUDA_CUDA(cudaMemset(d.counter, 0, 256));
    GUDA_CUDA(cudaHostAlloc(&d.host_scalar, 64, cudaHostAllocMapped));
    GUDA_CUDA(cudaHostGetDevicePointer(&d.host_scalar_dev, d.host_scalar, 0));
    it = g_dot_slots.emplace(s, d).first;
  }
  *out = it->second;
  return GUDA_OK;
}

template <typename T>
static int axpy_impl(int n, T alpha, const T* x, int incx, T* y, int incy, guda_stream_t stream, const char* op) {
  if (incx == 0) GUDA_PANIC("blas: zero x index increment");
  if (incy == 0) GUDA_PANIC("blas: zero y index increment");
  if (n < 1) {
    if (n == 0) return GUDA_OK;
    GUDA_PANIC("blas: n < 0");
  }
  int rc = bind_device();
  if (rc) return rc;
  if (!x || !y) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", op);
  if (alpha == T(0)) return GUDA_OK;                       // level1float32.go:243-245
  cudaStream_t s = native(stream);
  size_t N = (size_t)n;
  // fp64 rule: y's peel to 32-byte alignment and the last (n-peel)%4 elements are unfused
  size_t flo = 0, fhi = N;
  if (sizeof(T) == 8) {
    size_t peel = ((32 - ((uintptr_t)y & 31)) & 31) / 8;
    if (peel > N) peel = N;
    flo = peel;
    fhi = peel + ((N - peel) & ~(size_t)3);
  }
  if (incx == 1 && incy == 1) {
    if ((((uintptr_t)x ^ (uintptr_t)y) & 15) == 0 && ((uintptr_t)x % sizeof(T)) == 0) {
      size_t head = head_elems<T>(x);
      if (head > N) head = N;
      size_t nvec = (N - head) / Vec<T>::N;
      size_t chunk = (size_t)L1_THREADS * L1_UNROLL;
      unsigned grid = (unsigned)((nvec + chunk - 1) / chunk);
      if (grid == 0) grid = 1;
      axpy_unit_kernel<T><<<grid, L1_THREADS, 0, s>>>(N, alpha, x, y, head, flo, fhi);
    } else {
      unsigned grid = (unsigned)((N + L1_THREADS - 1) / L1_THREADS);
      if (grid > (unsigned)sm_count() * 16) grid = sm_count() * 16;
      axpy_strided_kernel<T, false><<<grid, L1_THREADS, 0, s>>>(N, alpha, x, 1, y, 1, flo, fhi);
    }
    return check_launch(op);
  }
  // strided: AxpyInc is unfused for both precisions (axpyinc_amd64.s); negative increments walk backwards
  const T* xs = x; T* ys = y;
  if (incx < 0) xs = x + (long long)(-n + 1) * incx;
  if (incy < 0) ys = y + (long long)(-n + 1) * incy;
  unsigned grid = (unsigned)((N + L1_THREADS - 1) / L1_THREADS);
  if (grid > (unsigned)sm_count() * 16) grid = sm_count() * 16;
  axpy_strided_kernel<T, false><<<grid, L1_THREADS, 0, s>>>(N, alpha, xs, incx, ys, incy, 0, 0);
  return check_launch(op);
}

template <typename T>
static int scal_impl(int n, T alpha, T* x, int incx, guda_stream_t stream, const char* op) {
  if (incx < 1) {
    if (incx == 0) GUDA_PANIC("blas: zero x index increment");
    return GUDA_OK;                                         // level1float32.go:625: no effect if incX < 0
  }
  if (n < 1) {
    if (n == 0) return GUDA_OK;
    GUDA_PANIC("blas: n < 0");
  }
  int rc = bind_device();
  if (rc) return rc;
  if (!x) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", op);
  cudaStream_t s = native(stream);
  size_t N = (size_t)n;
  int zero = (alpha == T(0));
  if (incx == 1 && ((uintptr_t)x % sizeof(T)) == 0) {
    size_t head = head_elems<T>(x);
    if (head > N) head = N;
    size_t nvec = (N - head) / Vec<T>::N;
    size_t chunk = (size_t)L1_THREADS * L1_UNROLL;
    unsigned grid = (unsigned)((nvec + chunk - 1) / chunk);
    if (grid == 0) grid = 1;
    scal_unit_kernel<T><<<grid, L1_THREADS, 0, s>>>(N, alpha, x, head, zero);
  } else {
    unsigned grid = (unsigned)((N + L1_THREADS - 1) / L1_THREADS);
    if (grid > (unsigned)sm_count() * 16) grid = sm_count() * 16;
    scal_strided_kernel<T><<<grid, L1_THREADS, 0, s>>>(N, alpha, x, incx, zero);
  }
  return check_launch(op);
}


template <typename T>
static int dot_impl(int n, const T* x, int incx, const T* y, int incy, T* host_result, T* dev_result, guda_stream_t stream, const char* op) {
  if (host_result) *host_result = T(0);
  if (incx == 0) GUDA_PANIC("blas: zero x index increment");
  if (incy == 0) GUDA_PANIC("blas: zero y index increment");
  int rc = bind_device();
  if (rc) return rc;
  cudaStream_t s = native(stream);
  if (n <= 0) {
    if (n < 0) GUDA_PANIC("blas: n < 0");
    if (dev_result) GUDA_CUDA(cudaMemsetAsync(dev_result, 0, sizeof(T), s));
    return GUDA_OK;                                         // level1float32_sdot.go:25-30: returns 0
  }
  if (!x || !y) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", op);
  size_t N = (size_t)n;
  const T* xs = x; const T* ys = y;
  if (incx < 0) xs = x + (long long)(-n + 1) * incx;
  if (incy < 0) ys = y + (long long)(-n + 1) * incy;
  bool unit = (incx == 1 && incy == 1);
  long long ix = incx, iy = incy;
  size_t head = 0;
  if (unit) {
    if ((((uintptr_t)x ^ (uintptr_t)y) & 15) == 0 && ((uintptr_t)x % sizeof(T)) == 0) {
      head = head_elems<T>(x);
      if (head > N) head = N;
    } else {
      unit = false;           // mutually misaligned: scalar path with stride 1
    }
  }
  size_t work = unit ? (N / Vec<T>::N + (size_t)L1_THREADS * L1_UNROLL - 1) / ((size_t)L1_THREADS * L1_UNROLL)
                     : (N + L1_THREADS - 1) / L1_THREADS;
  unsigned grid = (unsigned)work;
  static int occ = 0;       // resident CTAs per SM for this kernel: the grid must be exactly one full wave
  if (occ == 0) {
    int o = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, dot_kernel<T>, L1_THREADS, 0) != cudaSuccess || o < 1) o = 2;
    if (o > 4) o = 4;
    occ = o;
  }
  unsigned wave = (unsigned)sm_count() * (unsigned)occ;
  grid = (unsigned)dot_grid(work, wave);
  DotSlot slot;
  if ((rc = dot_slot_for(s, &slot))) return rc;
  T* partials = reinterpret_cast<T*>(slot.partials);                      // 8 KiB >= 592 partials of 8 bytes
  unsigned int* counter = slot.counter;
  T* host_alias = host_result ? reinterpret_cast<T*>(slot.host_scalar_dev) : nullptr;
  {
    cudaError_t le = launch_pdl(1, dot_kernel<T>, dim3(grid), dim3(L1_THREADS), 0, s, N, xs, (long long)(unit ? 1 : ix), ys, (long long)(unit ? 1 : iy),
                                head, partials, counter, host_alias, dev_result);
    if (le != cudaSuccess) return set_error(GUDA_ERR_EXECUTION, "%s launch failed: %s", op, cudaGetErrorString(le));
  }
  rc = check_launch(op);
  if (rc) return rc;
  if (host_result) {
    GUDA_CUDA(cudaStreamSynchronize(s));
    *host_result = *reinterpret_cast<volatile T*>(slot.host_scalar);
  }
  return GUDA_OK;
}

// ---------------------------------------------------------------- DOT with no allocation anywhere on the call path
// No counterpart in the reference: added for callers that cannot allocate (CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  Same kernel as guda_sdot_device; its scratch -- one fp32
// partial per CTA and the ticket counter -- lives in a file-scope __device__ array instead of the per-stream DotSlot that
// dot_slot_for() cudaMallocs on first use.  Every GPU that loads the library holds these 4352 bytes until exit.  In host code
// the array's NAME is a host-side stand-in, not a device address: the device address is resolved with cudaGetSymbolAddress, once
// per device, with that device current, and that pointer is what the kernel gets.  All calls on a device share the array: the
// ticket is cleared (stream-ordered) before every launch, partials are fully rewritten by every launch, and calls that OVERLAP on
// the GPU (different streams, or several host threads on different streams) race on it -- callers serialise them on one stream.
constexpr int DOT_STATIC_MAX_PARTIALS = 1024;                                        // grid is min(one wave, this)
constexpr size_t DOT_STATIC_COUNTER_OFF = DOT_STATIC_MAX_PARTIALS * sizeof(float);     // the ticket gets a 256-byte line of its own
constexpr size_t DOT_STATIC_BYTES = DOT_STATIC_COUNTER_OFF + 256;
static __device__ __align__(256) unsigned char g_dot_static_scratch[DOT_STATIC_BYTES];
constexpr int DOT_STATIC_MAX_DEVICES = 64;
static std::mutex g_dot_static_mu;
static unsigned char* g_dot_static_ptr[DOT_STATIC_MAX_DEVICES];                        // device address of the array, per device ordinal

static int dot_static_scratch(unsigned char** out) {
  int dev = -1;
  GUDA_CUDA(cudaGetDevice(&dev));                           // bind_device() has made the library's device current on this thread
  if (dev < 0 || dev >= DOT_STATIC_MAX_DEVICES) return set_error(GUDA_ERR_DEVICE, "Sdot (static scratch): device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_dot_static_mu);
  if (!g_dot_static_ptr[dev]) {
    unsigned char* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_dot_static_scratch);
    if (e != cudaSuccess || !ptr)
      return set_error(GUDA_ERR_DEVICE, "Sdot (static scratch): cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    g_dot_static_ptr[dev] = ptr;
  }
  *out = g_dot_static_ptr[dev];
  return GUDA_OK;
}

static int sdot_static_impl(int n, const float* x, int incx, const float* y, int incy, float* dev_result, guda_stream_t stream) {
  const char* op = "Sdot (static scratch)";
  if (incx == 0) GUDA_PANIC("blas: zero x index increment");
  if (incy == 0) GUDA_PANIC("blas: zero y index increment");
  if (n < 0) GUDA_PANIC("blas: n < 0");
  if (n > GUDA_SDOT_STATIC_MAX_N)
    return set_error(GUDA_ERR_INVALID_ARG, "%s: n = %d exceeds the maximum of %d elements this entry point accepts", op, n, GUDA_SDOT_STATIC_MAX_N);
  if (!runtime_inited())                                    // initialising allocates (runtime.cu init_locked); this call must not
    return set_error(GUDA_ERR_INVALID_ARG, "%s: library not initialised -- call guda_init first (this entry point allocates nothing, "
                                           "so it cannot set the library up itself)", op);
  int rc = bind_device();
  if (rc) return rc;
  cudaStream_t s = native(stream);
  if (n == 0) {
    GUDA_CUDA(cudaMemsetAsync(dev_result, 0, sizeof(float), s));          // level1float32_sdot.go:25-30: returns 0
    return GUDA_OK;
  }
  if (!x || !y) return set_error(GUDA_ERR_INVALID_ARG, "%s: null pointer", op);
  const size_t N = (size_t)n;
  const float* xs = x; const float* ys = y;
  if (incx < 0) xs = x + (long long)(-n + 1) * incx;
  if (incy < 0) ys = y + (long long)(-n + 1) * incy;
  bool unit = (incx == 1 && incy == 1);
  size_t head = 0;
  if (unit) {
    if ((((uintptr_t)x ^ (uintptr_t)y) & 15) == 0 && ((uintptr_t)x % sizeof(float)) == 0) {
      head = head_elems<float>(x);
      if (head > N) head = N;
    } else {
      unit = false;
    }
  }
  const size_t work = unit ? (N / Vec<float>::N + (size_t)L1_THREADS * L1_UNROLL - 1) / ((size_t)L1_THREADS * L1_UNROLL)
                           : (N + L1_THREADS - 1) / L1_THREADS;
  static int occ = 0;
  if (occ == 0) {
    int o = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, dot_kernel<float>, L1_THREADS, 0) != cudaSuccess || o < 1) o = 2;
    occ = o > 4 ? 4 : o;
  }
  size_t wave = (size_t)sm_count() * (size_t)occ;
  if (wave > (size_t)DOT_STATIC_MAX_PARTIALS) wave = DOT_STATIC_MAX_PARTIALS;
  const size_t grid = dot_grid(work, wave);
  unsigned char* scratch = nullptr;
  if ((rc = dot_static_scratch(&scratch))) return rc;
  float* partials = reinterpret_cast<float*>(scratch);
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch + DOT_STATIC_COUNTER_OFF);
  GUDA_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned int), s));        // the array keeps the last call's contents: the ticket must start at 0
  dot_kernel<float><<<(unsigned)grid, L1_THREADS, 0, s>>>(N, xs, (long long)(unit ? 1 : incx), ys, (long long)(unit ? 1 : incy), head, partials,
                                                           counter, (float*)nullptr, dev_result);
  return check_launch(op);
}

}  // namespace guda

using namespace guda;
extern "C" {
int guda_saxpy(int n, float alpha, const float* x, int incx, float* y, int incy, guda_stream_t stream) {
  return axpy_impl<float>(n, alpha, x, incx, y, incy, stream, "Saxpy");
}
int guda_daxpy(int n, double alpha, const double* x, int incx, double* y, int incy, guda_stream_t stream) {
  return axpy_impl<double>(n, alpha, x, incx, y, incy, stream, "Daxpy");
}
int guda_sscal(int n, float alpha, float* x, int incx, guda_stream_t stream) { return scal_impl<float>(n, alpha, x, incx, stream, "Sscal"); }
int guda_dscal(int n, double alpha, double* x, int incx, guda_stream_t stream) { return scal_impl<double>(n, alpha, x, incx, stream, "Dscal"); }
int guda_sdot(int n, const float* x, int incx, const float* y, int incy, float* host_result, guda_stream_t stream) {
  if (!host_result) return set_error(GUDA_ERR_INVALID_ARG, "Sdot: null result pointer");
  return dot_impl<float>(n, x, incx, y, incy, host_result, nullptr, stream, "Sdot");
}
int guda_sdot_device(int n, const float* x, int incx, const float* y, int incy, float* dev_result, guda_stream_t stream) {
  if (!dev_result) return set_error(GUDA_ERR_INVALID_ARG, "Sdot: null result pointer");
  return dot_impl<float>(n, x, incx, y, incy, nullptr, dev_result, stream, "Sdot");
}
int guda_ddot(int n, const double* x, int incx, const double* y, int incy, double* host_result, guda_stream_t stream) {
  if (!host_result) return set_error(GUDA_ERR_INVALID_ARG, "Ddot: null result pointer");
  return dot_impl<double>(n, x, incx, y, incy, host_result, nullptr, stream, "Ddot");
}
int guda_ddot_device(int n, const double* x, int incx, const double* y, int incy, double* dev_result, guda_stream_t stream) {
  if (!dev_result) return set_error(GUDA_ERR_INVALID_ARG, "Ddot: null result pointer");
  return dot_impl<double>(n, x, incx, y, incy, nullptr, dev_result, stream, "Ddot");
}
int guda_sdot_device_static(int n, const float* x, int incx, const float* y, int incy, float* dev_result, guda_stream_t stream) {
  if (!dev_result) return set_error(GUDA_ERR_INVALID_ARG, "Sdot (static scratch): null result pointer");
  return sdot_static_impl(n, x, incx, y, incy, dev_result, stream);
}
}
