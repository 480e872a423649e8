// guda.hpp -- C++ host-side mirror of GUDA's root-package API (package guda) above the C ABI of include/guda_b200.h.
//
// The reference is compiled Go; the build image has no Go toolchain, so besides the cgo shim (go/guda) this header gives
// compiled callers the same surface: same names, argument meaning and error behaviour.
//   Malloc/Free/Memcpy/DevicePtr      memory.go:57-66,81-145,246-317        AXPY/DOT/Scale/GEMM     math.go:39,64,170,98
//   Synchronize/CreateStream/Stream   guda.go:242-305                       FusedGEMMBias{ReLU,GELU} fused.go:313, fused_gelu.go:65
// Go `error` returns become exceptions: GUDAError (errors.go:27-44), ErrNotSupported (fused.go:11-17); Go panics with the
// fixed strings of compute/errors.go:9-35 become BlasPanic.  Header-only; link with -lguda_b200.
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "../include/guda_b200.h"

namespace guda {

enum class ErrorType { Memory = 0, InvalidArg, Execution, Numerical, Device, NotImplemented };   // errors.go:11-24

struct GUDAError : std::runtime_error {
  ErrorType type; std::string op, message;
  GUDAError(ErrorType t, std::string o, std::string m)
      : std::runtime_error("GUDA error in " + o + ": " + m), type(t), op(std::move(o)), message(std::move(m)) {}
};
struct ErrNotSupported : GUDAError { ErrNotSupported() : GUDAError(ErrorType::NotImplemented, "Fused", "operation not supported") {} };
struct ErrDoubleFree : GUDAError { ErrDoubleFree() : GUDAError(ErrorType::Memory, "Free", "double free detected") {} };
struct BlasPanic : std::logic_error { using std::logic_error::logic_error; };

inline void check(int rc, const char* op) {
  if (rc == GUDA_OK) return;
  std::string msg = guda_last_error();
  if (rc == GUDA_ERR_INVALID_ARG && msg.rfind("blas: ", 0) == 0) throw BlasPanic(msg);
  if (rc == GUDA_ERR_NOT_IMPLEMENTED && msg == "operation not supported") throw ErrNotSupported();
  if (rc == GUDA_ERR_MEMORY && msg == "double free detected") throw ErrDoubleFree();
  throw GUDAError(static_cast<ErrorType>(rc - 1), op, msg);
}

enum MemcpyKind { MemcpyHostToHost = 0, MemcpyHostToDevice, MemcpyDeviceToHost, MemcpyDeviceToDevice, MemcpyDefault };   // memory.go:15-21

// guda.go:95-99.  Typed views are device->host snapshots.
struct DevicePtr {
  void* ptr = nullptr; size_t size = 0; size_t offset = 0;
  DevicePtr Offset(size_t bytes) const { return DevicePtr{static_cast<char*>(ptr) + bytes, size - bytes, offset + bytes}; }
  size_t Size() const { return size; }
  template <typename T> std::vector<T> View() const {
    std::vector<T> out(size / sizeof(T));
    if (!out.empty()) check(guda_memcpy(out.data(), ptr, out.size() * sizeof(T), GUDA_MEMCPY_D2H, nullptr), "Memcpy");
    return out;
  }
  std::vector<float> Float32() const { return View<float>(); }
  std::vector<double> Float64() const { return View<double>(); }
  std::vector<int32_t> Int32() const { return View<int32_t>(); }
  std::vector<uint8_t> Byte() const { return View<uint8_t>(); }
  template <typename T> void CopyFrom(const std::vector<T>& src) const {
    if (src.size() * sizeof(T) > size) throw GUDAError(ErrorType::InvalidArg, "CopyFrom", "source larger than the device buffer");
    if (!src.empty()) check(guda_memcpy(ptr, src.data(), src.size() * sizeof(T), GUDA_MEMCPY_H2D, nullptr), "CopyFrom");
  }
};

inline DevicePtr Malloc(size_t size) { DevicePtr d; check(guda_malloc(&d.ptr, size), "Malloc"); d.size = size; return d; }
inline void Free(const DevicePtr& d) { check(guda_free(d.ptr), "Free"); }
inline void Memcpy(const DevicePtr& dst, const void* src, size_t size, MemcpyKind = MemcpyHostToDevice) {
  if (size) check(guda_memcpy(dst.ptr, src, size, GUDA_MEMCPY_H2D, nullptr), "Memcpy");
}
inline void Memcpy(void* dst, const DevicePtr& src, size_t size, MemcpyKind = MemcpyDeviceToHost) {
  if (size) check(guda_memcpy(dst, src.ptr, size, GUDA_MEMCPY_D2H, nullptr), "Memcpy");
}
inline void Memcpy(const DevicePtr& dst, const DevicePtr& src, size_t size, MemcpyKind = MemcpyDeviceToDevice) {
  if (size) check(guda_memcpy(dst.ptr, src.ptr, size, GUDA_MEMCPY_D2D, nullptr), "Memcpy");
}
inline void GetStats(int64_t* allocated, int64_t* peak) { check(guda_pool_stats(allocated, peak), "GetStats"); }   // memory.go:230
inline void Synchronize() { check(guda_device_synchronize(), "Synchronize"); }                                     // guda.go:200
inline int GetDeviceCount() { return 1; }                                                                         // guda.go:227-229
inline void SetDevice(int id) { if (id != 0) throw GUDAError(ErrorType::InvalidArg, "SetDevice", "invalid device ID"); }

class Stream {                                                                                                    // guda.go:57-62
 public:
  Stream() { check(guda_stream_create(&h_), "CreateStream"); }
  ~Stream() { if (h_) guda_stream_destroy(h_); }
  Stream(const Stream&) = delete; Stream& operator=(const Stream&) = delete;
  void Synchronize() { check(guda_stream_synchronize(h_), "StreamSynchronize"); }
  guda_stream_t handle() const { return h_; }
 private:
  guda_stream_t h_ = nullptr;
};

class Event {
 public:
  Event() { check(guda_event_create(&h_), "EventCreate"); }
  ~Event() { if (h_) guda_event_destroy(h_); }
  Event(const Event&) = delete; Event& operator=(const Event&) = delete;
  void Record(const Stream* s = nullptr) { check(guda_event_record(h_, s ? s->handle() : nullptr), "EventRecord"); }
  void Synchronize() { check(guda_event_synchronize(h_), "EventSynchronize"); }
  static float ElapsedTime(const Event& a, const Event& b) { float ms = 0; check(guda_event_elapsed_ms(a.h_, b.h_, &ms), "EventElapsedTime"); return ms; }
 private:
  guda_event_t h_ = nullptr;
};

inline float* f32(const DevicePtr& d) { return static_cast<float*>(d.ptr); }
inline double* f64(const DevicePtr& d) { return static_cast<double*>(d.ptr); }

// math.go:39,64,170
inline void AXPY(float alpha, const DevicePtr& x, const DevicePtr& y, int n) { check(guda_saxpy(n, alpha, f32(x), 1, f32(y), 1, nullptr), "AXPY"); }
inline float DOT(const DevicePtr& x, const DevicePtr& y, int n) { float r = 0; check(guda_sdot(n, f32(x), 1, f32(y), 1, &r, nullptr), "DOT"); return r; }
inline void Scale(float alpha, const DevicePtr& x, int n) { check(guda_sscal(n, alpha, f32(x), 1, nullptr), "Scale"); }

// compute/sgemm.go:116-136: the slice-length panics need the DevicePtr sizes, so they are raised here
inline void check_lengths(bool ta, bool tb, int m, int n, int k, int lda, int ldb, int ldc, size_t la, size_t lb, size_t lc) {
  if (m <= 0 || n <= 0 || k < 0) return;
  auto need = [](long long rows, long long ld, long long cols) { return (size_t)((rows - 1) * ld + cols); };
  if (la < (ta ? need(k, lda, m) : need(m, lda, k))) throw BlasPanic("blas: insufficient length of a");
  if (lb < (tb ? need(n, ldb, k) : need(k, ldb, n))) throw BlasPanic("blas: insufficient length of b");
  if (lc < need(m, ldc, n)) throw BlasPanic("blas: insufficient length of c");
}

// math.go:98
inline void GEMM(bool transA, bool transB, int m, int n, int k, float alpha, const DevicePtr& a, int lda, const DevicePtr& b, int ldb,
                 float beta, const DevicePtr& c, int ldc) {
  if (m >= 0 && n >= 0 && k >= 0 && lda >= 1 && ldb >= 1 && ldc >= 1) check_lengths(transA, transB, m, n, k, lda, ldb, ldc, a.size / 4, b.size / 4, c.size / 4);
  check(guda_sgemm(transA, transB, m, n, k, alpha, f32(a), lda, f32(b), ldb, beta, f32(c), ldc, nullptr), "GEMM");
}
inline void Dgemm(bool transA, bool transB, int m, int n, int k, double alpha, const DevicePtr& a, int lda, const DevicePtr& b, int ldb,
                  double beta, const DevicePtr& c, int ldc) {
  if (m >= 0 && n >= 0 && k >= 0 && lda >= 1 && ldb >= 1 && ldc >= 1) check_lengths(transA, transB, m, n, k, lda, ldb, ldc, a.size / 8, b.size / 8, c.size / 8);
  check(guda_dgemm(transA, transB, m, n, k, alpha, f64(a), lda, f64(b), ldb, beta, f64(c), ldc, nullptr), "Dgemm");
}
// fused.go:313, fused_gelu.go:65
inline void FusedGEMMBiasReLU(bool transA, bool transB, int m, int n, int k, float alpha, const DevicePtr& a, int lda, const DevicePtr& b,
                              int ldb, float beta, const DevicePtr& c, int ldc, const DevicePtr& bias) {
  check(guda_sgemm_bias_relu(transA, transB, m, n, k, alpha, f32(a), lda, f32(b), ldb, beta, f32(c), ldc, f32(bias), nullptr), "FusedGEMMBiasReLU");
}
inline void FusedGEMMBiasGELU(bool transA, bool transB, int m, int n, int k, float alpha, const DevicePtr& a, int lda, const DevicePtr& b,
                              int ldb, float beta, const DevicePtr& c, int ldc, const DevicePtr& bias) {
  check(guda_sgemm_bias_gelu(transA, transB, m, n, k, alpha, f32(a), lda, f32(b), ldb, beta, f32(c), ldc, f32(bias), nullptr), "FusedGEMMBiasGELU");
}
// bfloat16.go:104, float16.go:201
inline void GEMMBFloat16(bool transA, bool transB, int m, int n, int k, float alpha, const DevicePtr& a, int lda, const DevicePtr& b, int ldb,
                         float beta, const DevicePtr& c, int ldc) {
  check(guda_gemm_bf16(transA, transB, m, n, k, alpha, (const uint16_t*)a.ptr, lda, (const uint16_t*)b.ptr, ldb, beta, (uint16_t*)c.ptr, ldc, nullptr), "GEMMBFloat16");
}
inline void GEMMFloat16(bool transA, bool transB, int m, int n, int k, float alpha, const DevicePtr& a, int lda, const DevicePtr& b, int ldb,
                        float beta, const DevicePtr& c, int ldc) {
  check(guda_gemm_f16(transA, transB, m, n, k, alpha, (const uint16_t*)a.ptr, lda, (const uint16_t*)b.ptr, ldb, beta, (uint16_t*)c.ptr, ldc, nullptr), "GEMMFloat16");
}

// compute/level1float32_dsdot.go:18 / level1float32_sdsdot.go:18 (unit-stride length checks here; the rest in the C ABI)
inline double Dsdot(int n, const DevicePtr& x, int incX, const DevicePtr& y, int incY) {
  if (n > 0 && incX == 1 && incY == 1) {
    if (x.size / 4 < (size_t)n) throw BlasPanic("blas: insufficient length of x");
    if (y.size / 4 < (size_t)n) throw BlasPanic("blas: insufficient length of y");
  }
  double r = 0; check(guda_dsdot(n, f32(x), incX, f32(y), incY, &r, nullptr), "Dsdot"); return r;
}
inline float Sdsdot(int n, float alpha, const DevicePtr& x, int incX, const DevicePtr& y, int incY) {
  if (n > 0 && incX == 1 && incY == 1) {
    if (x.size / 4 < (size_t)n) throw BlasPanic("blas: insufficient length of x");
    if (y.size / 4 < (size_t)n) throw BlasPanic("blas: insufficient length of y");
  }
  float r = 0; check(guda_sdsdot(n, alpha, f32(x), incX, f32(y), incY, &r, nullptr), "Sdsdot"); return r;
}

// GEMM on host-resident operands: copies pipelined under the kernel (guda_sgemm_host); same arguments, panics and bits as GEMM
inline void GEMMHost(bool transA, bool transB, int m, int n, int k, float alpha, const float* a, size_t lenA, int lda, const float* b, size_t lenB,
                     int ldb, float beta, float* c, size_t lenC, int ldc) {
  if (m >= 0 && n >= 0 && k >= 0 && lda >= 1 && ldb >= 1 && ldc >= 1) check_lengths(transA, transB, m, n, k, lda, ldb, ldc, lenA, lenB, lenC);
  check(guda_sgemm_host(transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, nullptr), "GEMM");
}

// reduce.go:13-56 (methods of DevicePtr in the reference; free functions here) and Isamax (compute/level1float32.go:89-130)
inline void slice_check(const DevicePtr& d, long long n) { if (n < 0 || (size_t)n > d.size / 4) throw std::out_of_range("slice bounds out of range"); }
inline float Sum(const DevicePtr& d, long long n) { slice_check(d, n); float r = 0; check(guda_ssum(n, f32(d), &r, nullptr), "Sum"); return r; }
inline float SumSquares(const DevicePtr& d, long long n) { slice_check(d, n); float r = 0; check(guda_ssumsq(n, f32(d), &r, nullptr), "SumSquares"); return r; }
inline float Max(const DevicePtr& d, long long n) { slice_check(d, n); float r = 0; check(guda_smax(n, f32(d), &r, nullptr), "Max"); return r; }
inline float Min(const DevicePtr& d, long long n) { slice_check(d, n); float r = 0; check(guda_smin(n, f32(d), &r, nullptr), "Min"); return r; }
inline long long ArgMax(const DevicePtr& d, long long n) { slice_check(d, n); long long r = -1; check(guda_sargmax(n, f32(d), &r, nullptr), "ArgMax"); return r; }
inline long long ArgMin(const DevicePtr& d, long long n) { slice_check(d, n); long long r = -1; check(guda_sargmin(n, f32(d), &r, nullptr), "ArgMin"); return r; }
inline void Softmax(const DevicePtr& x, long long n) { slice_check(x, n); check(guda_softmax(f32(x), n, nullptr), "Softmax"); }   // reduce.go:180
inline void LayerNorm(const DevicePtr& x, const DevicePtr& gamma, const DevicePtr& beta, long long n, float eps) {            // fused.go:270
  slice_check(x, n);
  const bool affine = gamma.ptr && beta.ptr;
  check(guda_layer_norm(f32(x), affine ? f32(gamma) : nullptr, affine ? f32(beta) : nullptr, n, eps, nullptr), "LayerNorm");
}
inline long long Isamax(int n, const DevicePtr& x, int incX) {
  if (incX < 1) { if (incX == 0) throw BlasPanic("blas: zero x index increment"); return -1; }
  if ((long long)(x.size / 4) <= (long long)(n - 1) * incX) throw BlasPanic("blas: insufficient length of x");
  long long r = -1; check(guda_isamax(n, f32(x), incX, &r, nullptr), "Isamax"); return r;
}

// compute/level2float32.go:244-308 Sgemv (row-major); slice-length panics here, the rest from the C ABI with the same strings
inline void Sgemv(bool trans, int m, int n, float alpha, const DevicePtr& a, int lda, const DevicePtr& x, int incX, float beta,
                  const DevicePtr& y, int incY) {
  if (m > 0 && n > 0 && lda >= (n > 1 ? n : 1) && incX != 0 && incY != 0) {
    const long long lenX = trans ? m : n, lenY = trans ? n : m, lx = (long long)(x.size / 4), ly = (long long)(y.size / 4);
    if ((incX > 0 && (lenX - 1) * incX >= lx) || (incX < 0 && (1 - lenX) * incX >= lx)) throw BlasPanic("blas: insufficient length of x");
    if ((incY > 0 && (lenY - 1) * incY >= ly) || (incY < 0 && (1 - lenY) * incY >= ly)) throw BlasPanic("blas: insufficient length of y");
    if ((long long)(a.size / 4) < (long long)lda * (m - 1) + n) throw BlasPanic("blas: insufficient length of a");
  }
  check(guda_sgemv(trans, m, n, alpha, f32(a), lda, f32(x), incX, beta, f32(y), incY, nullptr), "Sgemv");
  // compute/asm/f32/gemv.go:61-64: GemvT with beta == 0 and incY == 1 clears the whole y slice
  if (trans && m > 0 && n > 0 && alpha != 0 && beta == 0 && incY == 1 && y.size / 4 > (size_t)n)
    check(guda_memset(static_cast<char*>(y.ptr) + 4 * (size_t)n, 0, (y.size / 4 - n) * 4, nullptr), "Sgemv");
}
// compute/level2float32.go:24-65 Sger
inline void Sger(int m, int n, float alpha, const DevicePtr& x, int incX, const DevicePtr& y, int incY, const DevicePtr& a, int lda) {
  if (m > 0 && n > 0 && lda >= (n > 1 ? n : 1) && incX != 0 && incY != 0) {
    const long long lx = (long long)(x.size / 4), ly = (long long)(y.size / 4);
    if ((incX > 0 && lx <= (long long)(m - 1) * incX) || (incX < 0 && lx <= (long long)(1 - m) * incX)) throw BlasPanic("blas: insufficient length of x");
    if ((incY > 0 && ly <= (long long)(n - 1) * incY) || (incY < 0 && ly <= (long long)(1 - n) * incY)) throw BlasPanic("blas: insufficient length of y");
    if ((long long)(a.size / 4) < (long long)lda * (m - 1) + n) throw BlasPanic("blas: insufficient length of a");
  }
  check(guda_sger(m, n, alpha, f32(x), incX, f32(y), incY, f32(a), lda, nullptr), "Sger");
}

// compute/level2float64.go:232-300 Dgemv (fp64 twin; gemvN_avx2_amd64.s / gemvT_amd64.s clear or scale exactly len(y) elements)
inline void Dgemv(bool trans, int m, int n, double alpha, const DevicePtr& a, int lda, const DevicePtr& x, int incX, double beta,
                  const DevicePtr& y, int incY) {
  if (m > 0 && n > 0 && lda >= (n > 1 ? n : 1) && incX != 0 && incY != 0) {
    const long long lenX = trans ? m : n, lenY = trans ? n : m, lx = (long long)(x.size / 8), ly = (long long)(y.size / 8);
    if ((incX > 0 && (lenX - 1) * incX >= lx) || (incX < 0 && (1 - lenX) * incX >= lx)) throw BlasPanic("blas: insufficient length of x");
    if ((incY > 0 && (lenY - 1) * incY >= ly) || (incY < 0 && (1 - lenY) * incY >= ly)) throw BlasPanic("blas: insufficient length of y");
    if ((long long)(a.size / 8) < (long long)lda * (m - 1) + n) throw BlasPanic("blas: insufficient length of a");
  }
  check(guda_dgemv(trans, m, n, alpha, f64(a), lda, f64(x), incX, beta, f64(y), incY, nullptr), "Dgemv");
}
// compute/level2float64.go:24-65 Dger
inline void Dger(int m, int n, double alpha, const DevicePtr& x, int incX, const DevicePtr& y, int incY, const DevicePtr& a, int lda) {
  if (m > 0 && n > 0 && lda >= (n > 1 ? n : 1) && incX != 0 && incY != 0) {
    const long long lx = (long long)(x.size / 8), ly = (long long)(y.size / 8);
    if ((incX > 0 && lx <= (long long)(m - 1) * incX) || (incX < 0 && lx <= (long long)(1 - m) * incX)) throw BlasPanic("blas: insufficient length of x");
    if ((incY > 0 && ly <= (long long)(n - 1) * incY) || (incY < 0 && ly <= (long long)(1 - n) * incY)) throw BlasPanic("blas: insufficient length of y");
    if ((long long)(a.size / 8) < (long long)lda * (m - 1) + n) throw BlasPanic("blas: insufficient length of a");
  }
  check(guda_dger(m, n, alpha, f64(x), incX, f64(y), incY, f64(a), lda, nullptr), "Dger");
}

// math.go:133-169, :230-272 elementwise device ops; fused.go:252 LinearReLU, :261 GELU (FusedKernel op lists)
inline void ew_len(const char* op, long long n, std::initializer_list<const DevicePtr*> ps) {
  for (const DevicePtr* p : ps) if (n < 0 || (long long)(p->size / 4) < n) throw std::out_of_range(std::string(op) + ": index out of range");
}
inline void Add(const DevicePtr& a, const DevicePtr& b, const DevicePtr& c, long long n) {
  ew_len("Add", n, {&a, &b, &c}); check(guda_add(f32(a), f32(b), f32(c), n, nullptr), "Add");
}
inline void Multiply(const DevicePtr& a, const DevicePtr& b, const DevicePtr& c, long long n) {
  ew_len("Multiply", n, {&a, &b, &c}); check(guda_multiply(f32(a), f32(b), f32(c), n, nullptr), "Multiply");
}
inline void ReLU(const DevicePtr& x, long long n) { ew_len("ReLU", n, {&x}); check(guda_relu(f32(x), n, nullptr), "ReLU"); }
inline void Sigmoid(const DevicePtr& x, long long n) { ew_len("Sigmoid", n, {&x}); check(guda_sigmoid(f32(x), n, nullptr), "Sigmoid"); }
inline void LinearReLU(const DevicePtr& x, float alpha, float beta, long long n) {
  ew_len("FusedOp", n, {&x});
  const guda_fused_op_t ops[3] = {{GUDA_FUSED_MUL_SCALAR, alpha, 0, 0}, {GUDA_FUSED_ADD_SCALAR, beta, 0, 0}, {GUDA_FUSED_RELU, 0.f, 0, 0}};
  check(guda_fused_execute(f32(x), nullptr, 0, n, ops, 3, nullptr), "FusedOp");
}
inline void GELU(const DevicePtr& x, long long n) {
  ew_len("FusedOp", n, {&x});
  const guda_fused_op_t op = {GUDA_FUSED_GELU, 0.f, 0, 0};
  check(guda_fused_execute(f32(x), nullptr, 0, n, &op, 1, nullptr), "FusedOp");
}

// reduce.go:105 ReduceSum / :139 ReduceMax over a row-major m x n matrix (axis 0 -> out[n], axis 1 -> out[m])
inline void ReduceSum(const DevicePtr& x, int m, int n, int axis, const DevicePtr& out) {
  ew_len("ReduceSum", (long long)m * n, {&x}); ew_len("ReduceSum", axis == 0 ? n : m, {&out});
  check(guda_reduce_sum_axis(f32(x), m, n, axis, f32(out), nullptr), "ReduceSum");
}
inline void ReduceMax(const DevicePtr& x, int m, int n, int axis, const DevicePtr& out) {
  ew_len("ReduceMax", (long long)m * n, {&x}); ew_len("ReduceMax", axis == 0 ? n : m, {&out});
  check(guda_reduce_max_axis(f32(x), m, n, axis, f32(out), nullptr), "ReduceMax");
}

// reduce.go:247 TopK: values (descending) and indices in the reference's selection-sort order; k clamped to n
inline void TopK(const DevicePtr& x, long long n, int k, std::vector<float>& values, std::vector<long long>& indices) {
  ew_len("TopK", n, {&x});
  const long long kk = k > n ? n : k;
  values.assign(kk > 0 ? (size_t)kk : 0, 0.f); indices.assign(kk > 0 ? (size_t)kk : 0, 0);
  check(guda_topk(f32(x), n, k, values.data(), indices.data(), nullptr), "TopK");
}

// conv.go:9-31 ConvParams, :69 Conv2D
struct ConvParams {
  int BatchSize = 0, InChannels = 0, InHeight = 0, InWidth = 0, OutChannels = 0, KernelHeight = 0, KernelWidth = 0;
  int StrideH = 1, StrideW = 1, PadH = 0, PadW = 0, DilationH = 1, DilationW = 1;
  bool UseBias = false;
  int OutputHeight() const { return (InHeight + 2 * PadH - ((KernelHeight - 1) * DilationH + 1)) / StrideH + 1; }
  int OutputWidth() const { return (InWidth + 2 * PadW - ((KernelWidth - 1) * DilationW + 1)) / StrideW + 1; }
};
inline void Conv2D(const DevicePtr& input, const DevicePtr& kernel, const DevicePtr& bias, const DevicePtr& output, const ConvParams& p) {
  guda_conv_params_t cp{p.BatchSize, p.InChannels, p.InHeight, p.InWidth, p.OutChannels, p.KernelHeight, p.KernelWidth,
                        p.StrideH, p.StrideW, p.PadH, p.PadW, p.DilationH, p.DilationW, p.UseBias ? 1 : 0};
  check(guda_conv2d(f32(input), f32(kernel), f32(bias), f32(output), &cp, nullptr), "Conv2D");
}
// conv.go:231 Conv2DDirect (the reference's direct loops, bit-identical; the cross-check API)
inline void Conv2DDirect(const DevicePtr& input, const DevicePtr& kernel, const DevicePtr& bias, const DevicePtr& output, const ConvParams& p) {
  guda_conv_params_t cp{p.BatchSize, p.InChannels, p.InHeight, p.InWidth, p.OutChannels, p.KernelHeight, p.KernelWidth,
                        p.StrideH, p.StrideW, p.PadH, p.PadW, p.DilationH, p.DilationW, p.UseBias ? 1 : 0};
  check(guda_conv2d_direct(f32(input), f32(kernel), f32(bias), f32(output), &cp, nullptr), "Conv2DDirect");
}

// ---- tolerance framework (tolerance.go:10-191, tolerance_arch.go:8-127, numerical_parity.go:124-141) + the sm_100a entry ----
struct ToleranceConfig {                                       // tolerance.go:10-26; the last four fields belong to the sm_100a entry
  float AbsTol = 0, RelTol = 0; int ULPTol = 0; bool CheckNaN = false, CheckInf = false;
  float KScale = 0; int HalfULPTol = 0; double Abs64 = 0, Rel64 = 0;
  guda_tolerance_t c() const { return guda_tolerance_t{AbsTol, RelTol, ULPTol, CheckNaN, CheckInf, KScale, HalfULPTol, Abs64, Rel64}; }
  static ToleranceConfig from(const guda_tolerance_t& t) {
    ToleranceConfig r; r.AbsTol = t.abs_tol; r.RelTol = t.rel_tol; r.ULPTol = t.ulp_tol; r.CheckNaN = t.check_nan != 0; r.CheckInf = t.check_inf != 0;
    r.KScale = t.k_scale; r.HalfULPTol = t.half_ulp_tol; r.Abs64 = t.abs64; r.Rel64 = t.rel64; return r;
  }
};
inline ToleranceConfig tolerance_preset(int which) { guda_tolerance_t t; check(guda_tolerance_preset(which, &t), "TolerancePreset"); return ToleranceConfig::from(t); }
inline ToleranceConfig DefaultTolerance() { return tolerance_preset(0); }      // tolerance.go:28-36
inline ToleranceConfig StrictTolerance() { return tolerance_preset(1); }       // tolerance.go:39-45
inline ToleranceConfig RelaxedTolerance() { return tolerance_preset(2); }      // tolerance.go:48-54
inline ToleranceConfig mergeTolerances(const ToleranceConfig& base, const ToleranceConfig& over) {                  // tolerance_arch.go:41-57
  guda_tolerance_t b = base.c(), o = over.c(), out; check(guda_merge_tolerances(&b, &o, &out), "mergeTolerances"); return ToleranceConfig::from(out);
}
// tolerance_arch.go:116-127; arch "" = the device that computed the results (sm_100a), or "amd64" / "arm64" / "generic"
inline ToleranceConfig GetOperationTolerance(const std::string& operation, const std::string& arch = "") {
  guda_tolerance_t t; check(guda_get_operation_tolerance(operation.c_str(), arch.empty() ? nullptr : arch.c_str(), &t), "GetOperationTolerance");
  return ToleranceConfig::from(t);
}
inline int Float32ULPDiff(float a, float b) {                                  // tolerance.go:100-118
  uint32_t x, y; std::memcpy(&x, &a, 4); std::memcpy(&y, &b, 4);
  if ((x ^ y) >> 31) return 0x7FFFFFFF;
  return static_cast<int>(x > y ? x - y : y - x);
}
inline bool Float32NearEqual(float a, float b, const ToleranceConfig& tol) {   // tolerance.go:57-97
  if (tol.CheckNaN && a != a && b != b) return true;
  const bool ia = a > 3.4028234e38f || a < -3.4028234e38f, ib = b > 3.4028234e38f || b < -3.4028234e38f;
  if (tol.CheckInf && ia && ib && ((a > 0) == (b > 0))) return true;
  if (a == b) return true;
  const float d32 = a - b;
  const double gap = d32 < 0 ? -(double)d32 : (double)d32;
  const double aa = a < 0 ? -(double)a : (double)a, bb = b < 0 ? -(double)b : (double)b;
  if (gap <= (double)tol.AbsTol || gap <= (aa > bb ? aa : bb) * (double)tol.RelTol) return true;
  return tol.ULPTol > 0 && Float32ULPDiff(a, b) <= tol.ULPTol;
}
struct VerificationResult {                                                    // tolerance.go:121-128
  float MaxAbsError = 0, MaxRelError = 0; int MaxULPError = 0; long long NumErrors = 0, TotalItems = 0, FirstError = -1;
  bool IsAcceptable(const ToleranceConfig& tol) const {                        // tolerance.go:177-182
    return NumErrors == 0 || (MaxAbsError <= tol.AbsTol && MaxRelError <= tol.RelTol && MaxULPError <= tol.ULPTol);
  }
};
// tolerance.go:131-174 over two device buffers of n floats (one streaming pass on the GPU)
inline VerificationResult VerifyFloat32Array(const DevicePtr& expected, const DevicePtr& actual, long long n, const ToleranceConfig& tol) {
  guda_tolerance_t t = tol.c(); guda_verify_result_t r;
  check(guda_verify_f32(static_cast<const float*>(expected.ptr), static_cast<const float*>(actual.ptr), n, &t, &r, nullptr), "VerifyFloat32Array");
  VerificationResult v; v.MaxAbsError = r.max_abs_error; v.MaxRelError = r.max_rel_error; v.MaxULPError = r.max_ulp_error;
  v.NumErrors = r.num_errors; v.TotalItems = r.total_items; v.FirstError = r.first_error; return v;
}

// ---- kernel-launch surface (guda.go:66-89,184-196,310-339): host closures cannot run on the device; the entry points exist and refuse ----
struct Dim3 { int X = 1, Y = 1, Z = 1; int Size() const { return X * Y * Z; } };
struct ThreadID {
  Dim3 BlockIdx, ThreadIdx, BlockDim, GridDim;
  int Global() const { return BlockIdx.X * BlockDim.X + ThreadIdx.X; }
  int GlobalX() const { return Global(); }
  int GlobalY() const { return BlockIdx.Y * BlockDim.Y + ThreadIdx.Y; }
  int GlobalZ() const { return BlockIdx.Z * BlockDim.Z + ThreadIdx.Z; }
};
template <typename Kernel, typename... Args>
inline void Launch(Kernel&&, const Dim3& grid, const Dim3& block, Args&&...) {
  if (grid.Size() <= 0 || block.Size() <= 0) throw GUDAError(ErrorType::InvalidArg, "Launch", "grid and block dimensions must be positive");
  throw GUDAError(ErrorType::NotImplemented, "Launch", "host closures cannot run on the device: use the device-kernel entry points "
                  "(Add/Multiply/ReLU/Sigmoid, FusedKernel, LinearReLU, GELU, GEMMBFloat16/GEMMFloat16, the reductions)");
}

}  // namespace guda
