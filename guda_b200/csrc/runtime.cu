// runtime.cu -- device, memory pool, copies, streams, events.
// Replaces memory.go:26-317 (MemoryPool/Malloc/Free/Memcpy) and guda.go:57-62,209-305
// (Device/Stream/Synchronize) with real device allocations, CUDA streams and events.
#include "common.cuh"
#include <map>
#include <mutex>
#include <vector>

namespace guda {

static thread_local char g_err[512];
char* last_error_buf() { return g_err; }
int set_error(int status, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return status;
}

int map_cuda_error(cudaError_t e) {
  switch (e) {
    case cudaErrorMemoryAllocation: return GUDA_ERR_MEMORY;
    case cudaErrorInvalidValue:
    case cudaErrorInvalidDevicePointer:
    case cudaErrorInvalidMemcpyDirection: return GUDA_ERR_INVALID_ARG;
    case cudaErrorNoDevice:
    case cudaErrorInvalidDevice:
    case cudaErrorInsufficientDriver:
    case cudaErrorDevicesUnavailable: return GUDA_ERR_DEVICE;
    default: return GUDA_ERR_EXECUTION;
  }
}

// ---------------------------------------------------------------- global state
// A freed block may still be read or written by work queued on the stream(s) it was last used on.  Each free records an
// event per such stream; a later owner either runs on that same stream (stream order protects it), makes its own stream
// wait for the event (internal temporaries: no host stall) or waits on the host (public Malloc: the caller may use any stream).
struct Pending { cudaStream_t stream; cudaEvent_t ev; };
struct Block { size_t size; bool used; std::vector<Pending> pending; };
struct State {
  std::mutex mu;
  bool inited = false;
  int device = 0;
  int sms = 0;
  Stream default_stream{nullptr};
  Scratch scratch{};
  // memory pool (memory.go:26-38)
  std::map<void*, Block> allocated;
  std::vector<void*> free_list;      // insertion order, first-fit scan like the reference
  std::vector<cudaStream_t> live_streams;   // default stream + every guda_stream_create'd stream
  std::vector<cudaEvent_t> event_cache;
  int64_t total_alloc = 0, peak_alloc = 0;
  int sgemm_mode = GUDA_SGEMM_EXACT;
};
static State& st() { static State s; return s; }

static int init_locked(State& s, int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return set_error(GUDA_ERR_DEVICE, "no CUDA device available (%s); libguda_b200 has no CPU fallback",
                     e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  if (device < 0 || device >= count) return set_error(GUDA_ERR_INVALID_ARG, "invalid device ID: %d", device);
  GUDA_CUDA(cudaSetDevice(device));
  cudaDeviceProp p;
  GUDA_CUDA(cudaGetDeviceProperties(&p, device));
  if (p.major < 10)
    return set_error(GUDA_ERR_DEVICE, "device %d is sm_%d%d; libguda_b200 is built for sm_100a only", device, p.major, p.minor);
  s.device = device;
  s.sms = p.multiProcessorCount;
  GUDA_CUDA(cudaStreamCreateWithFlags(&s.default_stream.s, cudaStreamNonBlocking));
  s.live_streams.push_back(s.default_stream.s);
  GUDA_CUDA(cudaMalloc(&s.scratch.dev_partials, 64 << 10));
  GUDA_CUDA(cudaMalloc((void**)&s.scratch.dev_counter, 256));
  GUDA_CUDA(cudaMemset(s.scratch.dev_counter, 0, 256));
  GUDA_CUDA(cudaHostAlloc(&s.scratch.host_scalar, 256, cudaHostAllocMapped));
  GUDA_CUDA(cudaHostGetDevicePointer(&s.scratch.host_scalar_dev, s.scratch.host_scalar, 0));
  s.inited = true;
  return GUDA_OK;
}

static int ensure_init() {
  State& s = st();
  if (s.inited) return GUDA_OK;
  std::lock_guard<std::mutex> lk(s.mu);
  if (s.inited) return GUDA_OK;
  int dev = 0;
  if (const char* e = getenv("GUDA_DEVICE")) dev = atoi(e);
  return init_locked(s, dev);
}

int bind_device() {
  int rc = ensure_init();
  if (rc) return rc;
  static thread_local int bound = -1;
  if (bound != st().device) {
    GUDA_CUDA(cudaSetDevice(st().device));
    bound = st().device;
  }
  return GUDA_OK;
}
int current_device() { return st().device; }
bool runtime_inited() { return st().inited; }
int sm_count() { return st().sms; }
cudaStream_t native(guda_stream_t h) { return h ? reinterpret_cast<Stream*>(h)->s : st().default_stream.s; }
int get_scratch(Scratch** out) { *out = &st().scratch; return GUDA_OK; }
int sgemm_mode() { return st().sgemm_mode; }

}  // namespace guda

using namespace guda;

extern "C" {

const char* guda_last_error(void) { return last_error_buf(); }
const char* guda_version(void) { return "guda-b200 0.1 (sm_100a)"; }

int guda_init(int device) {
  State& s = st();
  std::lock_guard<std::mutex> lk(s.mu);
  if (s.inited) {
    if (device != s.device) return set_error(GUDA_ERR_INVALID_ARG, "already initialised on device %d", s.device);
    return GUDA_OK;
  }
  return init_locked(s, device);
}

int guda_device_count(int* count) {
  if (!count) return set_error(GUDA_ERR_INVALID_ARG, "null pointer");
  cudaError_t e = cudaGetDeviceCount(count);
  if (e != cudaSuccess) { *count = 0; return set_error(GUDA_ERR_DEVICE, "cudaGetDeviceCount: %s", cudaGetErrorString(e)); }
  return GUDA_OK;
}

// guda.go:216-221 SetDevice: the reference accepts only its single device; here only the bound one
int guda_set_device(int device) {
  int rc = bind_device();
  if (rc) return rc;
  if (device != st().device) return set_error(GUDA_ERR_INVALID_ARG, "invalid device ID");
  return GUDA_OK;
}
int guda_get_device(int* device) {
  int rc = bind_device();
  if (rc) return rc;
  *device = st().device;
  return GUDA_OK;
}

int guda_device_properties(int device, guda_device_props_t* props) {
  if (!props) return set_error(GUDA_ERR_INVALID_ARG, "null pointer");
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count)
    return set_error(GUDA_ERR_INVALID_ARG, "invalid device ID: %d", device);   // guda.go:232-236
  cudaDeviceProp p;
  GUDA_CUDA(cudaGetDeviceProperties(&p, device));
  memset(props, 0, sizeof(*props));
  props->id = device;
  strncpy(props->name, p.name, sizeof(props->name) - 1);
  props->total_mem = p.totalGlobalMem;
  props->sm_count = p.multiProcessorCount;
  props->max_threads = p.multiProcessorCount * p.maxThreadsPerMultiProcessor;
  props->cc_major = p.major; props->cc_minor = p.minor;
  props->l2_bytes = p.l2CacheSize;
  return GUDA_OK;
}

int guda_device_synchronize(void) {
  int rc = bind_device();
  if (rc) return rc;
  GUDA_CUDA(cudaDeviceSynchronize());
  return GUDA_OK;
}

int guda_mem_get_info(size_t* free_bytes, size_t* total_bytes) {
  int rc = bind_device();
  if (rc) return rc;
  GUDA_CUDA(cudaMemGetInfo(free_bytes, total_bytes));
  return GUDA_OK;
}

// ---------------------------------------------------------------- memory pool
// memory.go:150-204 Allocate.  Accounting uses the reference's 64-byte rounding so that
// GetStats() returns the same integers for the same call sequence; the device blocks
// themselves come from cudaMalloc (256-byte aligned).
}  // extern "C"

namespace guda {

static cudaEvent_t take_event(State& s) {
  if (!s.event_cache.empty()) { cudaEvent_t e = s.event_cache.back(); s.event_cache.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return e;
}

// use_stream: the one stream the new owner will run on, or ANY_STREAM (public Malloc: unknown -> wait on the host)
static const cudaStream_t ANY_STREAM = reinterpret_cast<cudaStream_t>(~(uintptr_t)0);

static int pool_alloc_impl(void** dptr, size_t size, cudaStream_t use_stream) {
  if (!dptr) return set_error(GUDA_ERR_INVALID_ARG, "null pointer");
  *dptr = nullptr;
  int rc = bind_device();
  if (rc) return rc;
  State& s = st();
  const size_t alignment = 64;
  size_t aligned = (size + alignment - 1) & ~(alignment - 1);
  std::vector<Pending> waits;
  void* p = nullptr;
  {
    std::lock_guard<std::mutex> lk(s.mu);
    for (size_t i = 0; i < s.free_list.size(); i++) {
      Block& b = s.allocated[s.free_list[i]];
      if (b.size >= aligned) {
        p = s.free_list[i];
        s.free_list.erase(s.free_list.begin() + i);
        b.used = true;
        waits.swap(b.pending);
        s.total_alloc += (int64_t)b.size;
        if (s.total_alloc > s.peak_alloc) s.peak_alloc = s.total_alloc;
        break;
      }
    }
    if (!p) {
      cudaError_t e = cudaMalloc(&p, aligned ? aligned : alignment);
      if (e == cudaErrorMemoryAllocation && !s.free_list.empty()) {
        // out of device memory with cached blocks: release them (cudaFree waits for their queued work) and retry once
        cudaGetLastError();
        for (void* f : s.free_list) {
          for (Pending& w : s.allocated[f].pending) s.event_cache.push_back(w.ev);
          cudaFree(f);
          s.allocated.erase(f);
        }
        s.free_list.clear();
        e = cudaMalloc(&p, aligned ? aligned : alignment);
      }
      if (e != cudaSuccess) {
        cudaGetLastError();
        return set_error(e == cudaErrorMemoryAllocation ? GUDA_ERR_MEMORY : map_cuda_error(e), "out of memory (%zu bytes): %s",
                         aligned, cudaGetErrorString(e));     // errors.go:113 ErrOutOfMemory
      }
      s.allocated[p] = Block{aligned, true, {}};
      s.total_alloc += (int64_t)aligned;
      if (s.total_alloc > s.peak_alloc) s.peak_alloc = s.total_alloc;
    }
  }
  // outside the lock: order the new owner after the previous owner's queued work
  for (Pending& w : waits) {
    if (use_stream == ANY_STREAM) cudaEventSynchronize(w.ev);
    else if (w.stream != use_stream) cudaStreamWaitEvent(use_stream, w.ev, 0);
  }
  if (!waits.empty()) {
    std::lock_guard<std::mutex> lk(s.mu);
    for (Pending& w : waits) s.event_cache.push_back(w.ev);
  }
  *dptr = p;
  return GUDA_OK;
}

// only_stream: the block was used on this stream alone (internal temporaries); ANY_STREAM: record on every live stream
static int pool_free_impl(void* dptr, cudaStream_t only_stream) {
  int rc = bind_device();
  if (rc) return rc;
  State& s = st();
  std::lock_guard<std::mutex> lk(s.mu);
  auto it = s.allocated.find(dptr);
  if (it == s.allocated.end()) return set_error(GUDA_ERR_MEMORY, "pointer not found in allocation pool");
  if (!it->second.used) return set_error(GUDA_ERR_MEMORY, "double free detected");    // errors.go:116 ErrDoubleFree
  it->second.used = false;
  it->second.pending.clear();
  auto record = [&](cudaStream_t q) {
    if (cudaStreamQuery(q) == cudaSuccess) return;          // idle stream: nothing can still touch the block
    cudaGetLastError();
    cudaEvent_t e = take_event(s);
    if (!e || cudaEventRecord(e, q) != cudaSuccess) { cudaGetLastError(); cudaStreamSynchronize(q); if (e) s.event_cache.push_back(e); return; }
    it->second.pending.push_back(Pending{q, e});
  };
  if (only_stream == ANY_STREAM) { for (cudaStream_t q : s.live_streams) record(q); }
  else record(only_stream);
  s.free_list.push_back(dptr);
  s.total_alloc -= (int64_t)it->second.size;
  return GUDA_OK;
}

int pool_alloc(void** dptr, size_t size, cudaStream_t use_stream) { return pool_alloc_impl(dptr, size, use_stream); }
int pool_free(void* dptr, cudaStream_t used_on) { return pool_free_impl(dptr, used_on); }

}  // namespace guda

extern "C" {

int guda_malloc(void** dptr, size_t size) { return pool_alloc_impl(dptr, size, ANY_STREAM); }

// memory.go:207-227 Free
int guda_free(void* dptr) { return pool_free_impl(dptr, ANY_STREAM); }

int guda_pool_stats(int64_t* allocated, int64_t* peak) {
  State& s = st();
  std::lock_guard<std::mutex> lk(s.mu);
  if (allocated) *allocated = s.total_alloc;
  if (peak) *peak = s.peak_alloc;
  return GUDA_OK;
}

int guda_pool_trim(void) {
  int rc = bind_device();
  if (rc) return rc;
  State& s = st();
  std::lock_guard<std::mutex> lk(s.mu);
  GUDA_CUDA(cudaDeviceSynchronize());
  for (void* f : s.free_list) {
    for (Pending& w : s.allocated[f].pending) s.event_cache.push_back(w.ev);
    cudaFree(f);
    s.allocated.erase(f);
  }
  s.free_list.clear();
  return GUDA_OK;
}

int guda_malloc_host(void** host_ptr, size_t size) {
  int rc = bind_device();
  if (rc) return rc;
  GUDA_CUDA(cudaHostAlloc(host_ptr, size ? size : 1, cudaHostAllocDefault));
  return GUDA_OK;
}
int guda_free_host(void* host_ptr) {
  int rc = bind_device();
  if (rc) return rc;
  GUDA_CUDA(cudaFreeHost(host_ptr));
  return GUDA_OK;
}

// memory.go:81-145 Memcpy: the reference ignores `kind`; we validate its range and let UVA
// infer the direction, so every call site that passed a sloppy kind keeps working.
static int memcpy_impl(void* dst, const void* src, size_t size, int kind, guda_stream_t stream, bool block) {
  int rc = bind_device();
  if (rc) return rc;
  if (kind < GUDA_MEMCPY_H2H || kind > GUDA_MEMCPY_DEFAULT) return set_error(GUDA_ERR_INVALID_ARG, "invalid memcpy kind %d", kind);
  if (!dst || !src || size == 0) return GUDA_OK;          // memory.go:140: nil or empty -> no-op
  cudaStream_t s = native(stream);
  GUDA_CUDA(cudaMemcpyAsync(dst, src, size, cudaMemcpyDefault, s));
  if (block) GUDA_CUDA(cudaStreamSynchronize(s));
  return GUDA_OK;
}
int guda_memcpy(void* dst, const void* src, size_t size, int kind, guda_stream_t stream) {
  return memcpy_impl(dst, src, size, kind, stream, true);
}
int guda_memcpy_async(void* dst, const void* src, size_t size, int kind, guda_stream_t stream) {
  return memcpy_impl(dst, src, size, kind, stream, false);
}
int guda_memset(void* dptr, int value, size_t size, guda_stream_t stream) {
  int rc = bind_device();
  if (rc) return rc;
  if (!dptr || size == 0) return GUDA_OK;
  GUDA_CUDA(cudaMemsetAsync(dptr, value, size, native(stream)));
  return GUDA_OK;
}

// ---------------------------------------------------------------- streams / events
int guda_stream_create(guda_stream_t* stream) {
  int rc = bind_device();
  if (rc) return rc;
  Stream* s = new Stream{nullptr};
  cudaError_t e = cudaStreamCreateWithFlags(&s->s, cudaStreamNonBlocking);
  if (e != cudaSuccess) { delete s; return set_error(map_cuda_error(e), "cudaStreamCreate: %s", cudaGetErrorString(e)); }
  {
    std::lock_guard<std::mutex> lk(st().mu);
    st().live_streams.push_back(s->s);
  }
  *stream = reinterpret_cast<guda_stream_t>(s);
  return GUDA_OK;
}
int guda_stream_destroy(guda_stream_t stream) {
  if (!stream) return set_error(GUDA_ERR_INVALID_ARG, "cannot destroy the default stream");
  int rc = bind_device();
  if (rc) return rc;
  Stream* s = reinterpret_cast<Stream*>(stream);
  GUDA_CUDA(cudaStreamSynchronize(s->s));
  {
    std::lock_guard<std::mutex> lk(st().mu);
    auto& v = st().live_streams;
    for (size_t i = 0; i < v.size(); i++) if (v[i] == s->s) { v.erase(v.begin() + i); break; }
  }
  GUDA_CUDA(cudaStreamDestroy(s->s));
  delete s;
  return GUDA_OK;
}
int guda_stream_synchronize(guda_stream_t stream) {
  int rc = bind_device();
  if (rc) return rc;
  GUDA_CUDA(cudaStreamSynchronize(native(stream)));
  return GUDA_OK;
}
int guda_stream_query(guda_stream_t stream, int* done) {
  int rc = bind_device();
  if (rc) return rc;
  cudaError_t e = cudaStreamQuery(native(stream));
  if (e == cudaSuccess) { *done = 1; return GUDA_OK; }
  if (e == cudaErrorNotReady) { *done = 0; return GUDA_OK; }
  return set_error(map_cuda_error(e), "cudaStreamQuery: %s", cudaGetErrorString(e));
}
int guda_stream_native(guda_stream_t stream, void** cuda_stream) {
  int rc = bind_device();
  if (rc) return rc;
  *cuda_stream = (void*)native(stream);
  return GUDA_OK;
}
int guda_stream_wait_event(guda_stream_t stream, guda_event_t event) {
  int rc = bind_device();
  if (rc) return rc;
  if (!event) return set_error(GUDA_ERR_INVALID_ARG, "null event");
  GUDA_CUDA(cudaStreamWaitEvent(native(stream), reinterpret_cast<Event*>(event)->e, 0));
  return GUDA_OK;
}
int guda_event_create(guda_event_t* event) {
  int rc = bind_device();
  if (rc) return rc;
  Event* ev = new Event{nullptr};
  cudaError_t e = cudaEventCreate(&ev->e);
  if (e != cudaSuccess) { delete ev; return set_error(map_cuda_error(e), "cudaEventCreate: %s", cudaGetErrorString(e)); }
  *event = reinterpret_cast<guda_event_t>(ev);
  return GUDA_OK;
}
int guda_event_destroy(guda_event_t event) {
  if (!event) return set_error(GUDA_ERR_INVALID_ARG, "null event");
  int rc = bind_device();
  if (rc) return rc;
  Event* ev = reinterpret_cast<Event*>(event);
  GUDA_CUDA(cudaEventDestroy(ev->e));
  delete ev;
  return GUDA_OK;
}
int guda_event_record(guda_event_t event, guda_stream_t stream) {
  int rc = bind_device();
  if (rc) return rc;
  if (!event) return set_error(GUDA_ERR_INVALID_ARG, "null event");
  GUDA_CUDA(cudaEventRecord(reinterpret_cast<Event*>(event)->e, native(stream)));
  return GUDA_OK;
}
int guda_event_synchronize(guda_event_t event) {
  int rc = bind_device();
  if (rc) return rc;
  if (!event) return set_error(GUDA_ERR_INVALID_ARG, "null event");
  GUDA_CUDA(cudaEventSynchronize(reinterpret_cast<Event*>(event)->e));
  return GUDA_OK;
}
int guda_event_query(guda_event_t event, int* done) {
  int rc = bind_device();
  if (rc) return rc;
  if (!event) return set_error(GUDA_ERR_INVALID_ARG, "null event");
  cudaError_t e = cudaEventQuery(reinterpret_cast<Event*>(event)->e);
  if (e == cudaSuccess) { *done = 1; return GUDA_OK; }
  if (e == cudaErrorNotReady) { *done = 0; return GUDA_OK; }
  return set_error(map_cuda_error(e), "cudaEventQuery: %s", cudaGetErrorString(e));
}
int guda_event_elapsed_ms(guda_event_t start, guda_event_t stop, float* ms) {
  int rc = bind_device();
  if (rc) return rc;
  if (!start || !stop || !ms) return set_error(GUDA_ERR_INVALID_ARG, "null event");
  GUDA_CUDA(cudaEventElapsedTime(ms, reinterpret_cast<Event*>(start)->e, reinterpret_cast<Event*>(stop)->e));
  return GUDA_OK;
}

int guda_set_sgemm_mode(int mode) {
  if (mode != GUDA_SGEMM_EXACT && mode != GUDA_SGEMM_3XTF32) return set_error(GUDA_ERR_INVALID_ARG, "unknown sgemm mode %d", mode);
  st().sgemm_mode = mode;
  return GUDA_OK;
}
int guda_get_sgemm_mode(int* mode) { *mode = st().sgemm_mode; return GUDA_OK; }

}  // extern "C"
