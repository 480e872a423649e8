// common.cuh -- shared plumbing for libguda_b200: status/last-error, stream handles, launch checks.
#pragma once
#include <cuda_runtime.h>
#include <vector>
#include <cstdarg>
#include <cstdlib>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include "../../include/guda_b200.h"

namespace guda {

// thread-local last error (guda_last_error)
char* last_error_buf();
int set_error(int status, const char* fmt, ...);

// Go panic strings (compute/errors.go:9-35), returned through guda_last_error with GUDA_ERR_INVALID_ARG
#define GUDA_PANIC(msg) return ::guda::set_error(GUDA_ERR_INVALID_ARG, "%s", msg)

#define GUDA_CUDA(call)                                                                              \
  do {                                                                                               \
    cudaError_t _e = (call);                                                                         \
    if (_e != cudaSuccess)                                                                           \
      return ::guda::set_error(::guda::map_cuda_error(_e), "%s failed: %s", #call, cudaGetErrorString(_e)); \
  } while (0)

int map_cuda_error(cudaError_t e);

// binds the library's device to the calling OS thread (cgo may migrate goroutines between threads)
int bind_device();
int current_device();
int sm_count();
bool runtime_inited();       // true once guda_init (or the first ordinary call) has set the library up; never initialises

struct Stream { cudaStream_t s; };
struct Event { cudaEvent_t e; };

// NULL handle = the library's default (non-blocking) stream
cudaStream_t native(guda_stream_t h);

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(GUDA_ERR_EXECUTION, "%s launch failed: %s", what, cudaGetErrorString(e));
  return GUDA_OK;
}

// Programmatic dependent launch (PDL): the kernel may be SCHEDULED while the previous kernel on the stream is still draining, and
// blocks at pdl_wait() until that kernel has completed and its writes are visible; pdl_trigger() tells the scheduler that the
// next kernel's CTAs may start arriving.  Hides the launch ramp between back-to-back kernels (a 20-50 us kernel pays ~2 us of
// it otherwise).  Both device calls are no-ops when the launch did not carry the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// which launch sites carry the attribute: bit 0 DOT, bit 1 SGEMM pre-pass, bit 2 SGEMM pipeline (GUDA_PDL_MASK, A/B knob)
inline int pdl_mask() {
  static const int m = [] { const char* e = getenv("GUDA_PDL_MASK"); return e ? atoi(e) : 7; }();
  return m;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int site_bit, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr{};
  attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &attr; cfg.numAttrs = (pdl_mask() & site_bit) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// scratch that lives as long as the library: pinned host scalars + device partials for reductions
struct Scratch {
  void* dev_partials;      // >= 64 KiB
  unsigned int* dev_counter;
  void* host_scalar;       // pinned, mapped
  void* host_scalar_dev;   // device alias of host_scalar
};
int get_scratch(Scratch** out);

// stream-aware pool entry points for internal temporaries (runtime.cu): the block is handed out for use on `use_stream`
// (a block last used on another stream makes use_stream wait for that work on the device -- no host stall) and returned
// with pool_free(ptr, stream) while its last kernels may still be queued on that stream.
int pool_alloc(void** dptr, size_t size, cudaStream_t use_stream);
int pool_free(void* dptr, cudaStream_t used_on);
struct PoolTemp {
  void* p = nullptr;
  cudaStream_t s = nullptr;
  PoolTemp() = default;
  PoolTemp(const PoolTemp&) = delete;
  PoolTemp& operator=(const PoolTemp&) = delete;
  ~PoolTemp() { if (p) pool_free(p, s); }
  int alloc(size_t bytes, cudaStream_t stream) { s = stream; return pool_alloc(&p, bytes, stream); }
};

// two library-lifetime non-blocking streams for copies pipelined under compute (host_gemm.cu)
int aux_copy_streams(cudaStream_t* in, cudaStream_t* out, cudaStream_t* in2 = nullptr);
// K boundaries of the host-operand pipeline (short first chunk, then 8192-deep)
void host_k_chunks(int k, std::vector<int>& bounds);
int host_c_panels(int m);

}  // namespace guda
