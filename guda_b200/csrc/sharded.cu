// sharded.cu -- multi-GPU forms (new, additive API): 2-D block-partitioned SGEMM + all-gather of C,
// sharded DOT + all-reduce.  One process per GPU; NCCL over NVLink 5 / NVSwitch, loaded at run time
// (dlopen) so the library has no link-time NCCL dependency and single-GPU users never touch it.
// The reference has no multi-device notion at all (guda.go:227-229 GetDeviceCount() == 1).
#include "common.cuh"
#include <string>
#include <chrono>
#include <cuda.h>
#include <dlfcn.h>
#include <nccl.h>
#include <mutex>
#include <vector>

namespace guda {

int sgemm_mode();
constexpr int GUDA_MAX_SCATTER_RANKS = 8;     // == GUDA_MAX_SCATTER of gemm_simt.cuh

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*CommSplit)(ncclComm_t, int, int, ncclComm_t*, void*) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;
static std::mutex g_nccl_mu;

static int load_nccl() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.handle) return GUDA_OK;
  // prefer a copy that is already in the process (e.g. the one torch loaded), then GUDA_NCCL_LIB, then the system one
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!h) { if (const char* e = getenv("GUDA_NCCL_LIB")) h = dlopen(e, RTLD_NOW | RTLD_GLOBAL); }
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return set_error(GUDA_ERR_DEVICE, "cannot load libnccl.so.2: %s", dlerror());
#define SYM(field, name)                                                                    \
  *(void**)(&g_nccl.field) = dlsym(h, name);                                                \
  if (!g_nccl.field) return set_error(GUDA_ERR_DEVICE, "libnccl is missing %s", name);
  SYM(GetUniqueId, "ncclGetUniqueId")
  SYM(CommInitRank, "ncclCommInitRank")
  SYM(CommDestroy, "ncclCommDestroy")
  SYM(AllGather, "ncclAllGather")
  SYM(AllReduce, "ncclAllReduce")
  SYM(Broadcast, "ncclBroadcast")
  SYM(CommSplit, "ncclCommSplit")
  SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
  g_nccl.handle = h;
  return GUDA_OK;
}

#define GUDA_NCCL(call)                                                                                      \
  do {                                                                                                       \
    ncclResult_t _r = (call);                                                                                \
    if (_r != ncclSuccess) return set_error(GUDA_ERR_EXECUTION, "%s failed: %s", #call, g_nccl.GetErrorString(_r)); \
  } while (0)

// what a rank publishes about its gathered-C buffer: the IPC handle of the allocation that holds it + the offset inside it
struct PeerRecord {
  cudaIpcMemHandle_t handle;     // 64 bytes
  unsigned long long offset;
  unsigned long long ok;
};
struct PeerMapping { cudaIpcMemHandle_t handle; void* base; };

struct Comm {
  ncclComm_t comm;
  int nranks, rank;
  void* staging = nullptr;       // all-gather landing zone (NCCL mode), grown on demand
  size_t staging_bytes = 0;
  float* dev_scalar = nullptr;
  int gather_mode = GUDA_GATHER_P2P;
  bool p2p_failed = false;       // agreed by all ranks: peer mapping is not available here
  PeerRecord* dev_records = nullptr;            // [nranks + 1]: slot nranks is this rank's send buffer
  PeerRecord* host_records = nullptr;           // pinned, [nranks + 1]
  std::vector<std::vector<PeerMapping>> opened; // per peer rank: allocations already mapped into this process
  // process-grid sub-communicators of the host-operand step (created on first use, collectively): ranks of one grid row share an
  // A row panel, ranks of one grid column a B column panel
  ncclComm_t row_comm = nullptr, col_comm = nullptr;
  int grid_pr = 0, grid_pc = 0;
  // steady state: the gathered-C buffers of the previous call, valid while no rank's (handle, offset) record changes
  bool cache_valid = false;
  PeerRecord cache_mine{};
  void* cache_peers[GUDA_MAX_SCATTER_RANKS] = {};
};

// Collective: every rank publishes where its c_full lives; returns device-usable pointers to every rank's c_full.
// The NCCL all-gather + stream synchronise also acts as the entry barrier: when it returns, every rank has finished
// whatever preceded this call on its stream, so peers may overwrite its c_full.
static int resolve_peers(Comm* c, void* c_full, cudaStream_t s, void** out, bool* all_ok) {
  PeerRecord mine{};
  CUdeviceptr base = 0; size_t range = 0;
  // the driver entry point is resolved at run time so that the library has no link-time dependency on libcuda.so.1
  typedef CUresult (*AddressRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
  static AddressRangeFn address_range = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) fn = nullptr;
    return reinterpret_cast<AddressRangeFn>(fn);
  }();
  bool ok = address_range && address_range(&base, &range, (CUdeviceptr)c_full) == CUDA_SUCCESS &&
            cudaIpcGetMemHandle(&mine.handle, (void*)base) == cudaSuccess;
  cudaGetLastError();
  mine.offset = ok ? (unsigned long long)((CUdeviceptr)c_full - base) : 0;
  mine.ok = ok ? 1 : 0;
  // Steady state (every step of a loop gathers into the same buffers): one all-reduce asks "did any rank's record change?".
  // It is also the entry barrier -- it completes on this stream only after every rank has reached it on theirs, i.e. after
  // whatever they queued before this call, so peers may now overwrite this rank's c_full.  Only a change pays the full exchange.
  float changed = (!c->cache_valid || memcmp(&mine, &c->cache_mine, sizeof(mine)) != 0) ? 1.0f : 0.0f;
  GUDA_CUDA(cudaMemcpyAsync(c->dev_scalar, &changed, sizeof(float), cudaMemcpyHostToDevice, s));
  GUDA_NCCL(g_nccl.AllReduce(c->dev_scalar, c->dev_scalar, 1, ncclFloat, ncclMax, c->comm, s));
  GUDA_CUDA(cudaMemcpyAsync(&changed, c->dev_scalar, sizeof(float), cudaMemcpyDeviceToHost, s));
  GUDA_CUDA(cudaStreamSynchronize(s));
  if (changed < 0.5f) {
    for (int r = 0; r < c->nranks; r++) out[r] = c->cache_peers[r];
    *all_ok = true;
    return GUDA_OK;
  }
  c->cache_valid = false;
  c->host_records[c->nranks] = mine;
  GUDA_CUDA(cudaMemcpyAsync(c->dev_records + c->nranks, c->host_records + c->nranks, sizeof(PeerRecord), cudaMemcpyHostToDevice, s));
  GUDA_NCCL(g_nccl.AllGather(c->dev_records + c->nranks, c->dev_records, sizeof(PeerRecord), ncclChar, c->comm, s));
  GUDA_CUDA(cudaMemcpyAsync(c->host_records, c->dev_records, sizeof(PeerRecord) * c->nranks, cudaMemcpyDeviceToHost, s));
  GUDA_CUDA(cudaStreamSynchronize(s));
  bool good = true;
  for (int r = 0; r < c->nranks; r++) good = good && c->host_records[r].ok;
  for (int r = 0; r < c->nranks && good; r++) {
    if (r == c->rank) { out[r] = c_full; continue; }
    const PeerRecord& rec = c->host_records[r];
    void* mapped = nullptr;
    for (const PeerMapping& pm : c->opened[r])
      if (memcmp(&pm.handle, &rec.handle, sizeof(cudaIpcMemHandle_t)) == 0) { mapped = pm.base; break; }
    if (!mapped) {
      if (cudaIpcOpenMemHandle(&mapped, rec.handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); good = false; break; }
      c->opened[r].push_back(PeerMapping{rec.handle, mapped});
    }
    out[r] = static_cast<char*>(mapped) + rec.offset;
  }
  // every rank must take the same route: agree on the outcome (a rank that could not map a peer vetoes P2P for all)
  float flag = good ? 1.0f : 0.0f;
  GUDA_CUDA(cudaMemcpyAsync(c->dev_scalar, &flag, sizeof(float), cudaMemcpyHostToDevice, s));
  GUDA_NCCL(g_nccl.AllReduce(c->dev_scalar, c->dev_scalar, 1, ncclFloat, ncclMin, c->comm, s));
  GUDA_CUDA(cudaMemcpyAsync(&flag, c->dev_scalar, sizeof(float), cudaMemcpyDeviceToHost, s));
  GUDA_CUDA(cudaStreamSynchronize(s));
  *all_ok = flag > 0.5f;
  if (*all_ok) {
    c->cache_valid = true;
    c->cache_mine = mine;
    for (int r = 0; r < c->nranks; r++) c->cache_peers[r] = out[r];
  }
  return GUDA_OK;
}

// gathered [rank][m_loc][n_loc] blocks -> row-major c_full; 128-bit copies when the geometry allows
template <typename T, typename V>
__global__ void unpack_blocks_kernel(const T* __restrict__ staging, T* __restrict__ c_full, long long ldc_full, int pr, int pc,
                                     int m_loc, int n_loc) {
  constexpr int VN = sizeof(V) / sizeof(T);
  const long long per_block = (long long)m_loc * n_loc;
  const long long total = per_block * pr * pc;
  if ((n_loc % VN) == 0 && (ldc_full % VN) == 0 && ((uintptr_t)c_full & 15) == 0 && ((uintptr_t)staging & 15) == 0) {
    const long long nv = total / VN;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += (long long)gridDim.x * blockDim.x) {
      long long e = v * VN;
      int q = (int)(e / per_block);
      long long w = e % per_block;
      int i = (int)(w / n_loc), j = (int)(w % n_loc);
      int r = q / pc, c = q % pc;
      V val = *reinterpret_cast<const V*>(staging + e);
      *reinterpret_cast<V*>(c_full + ((long long)r * m_loc + i) * ldc_full + (long long)c * n_loc + j) = val;
    }
  } else {
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
      int q = (int)(e / per_block);
      long long w = e % per_block;
      int i = (int)(w / n_loc), j = (int)(w % n_loc);
      int r = q / pc, c = q % pc;
      c_full[((long long)r * m_loc + i) * ldc_full + (long long)c * n_loc + j] = staging[e];
    }
  }
}

// local block -> every rank's c_full through NCCL (the route beside the fused epilogue)
template <typename T, typename V>
static int gather_blocks_nccl(Comm* c, int pr, int pc, int m_loc, int n_loc, int n_total, const T* c_block, T* c_full, int ldc_full,
                              ncclDataType_t dt, cudaStream_t s) {
  const size_t per_block = (size_t)m_loc * n_loc;
  if (pc == 1 && ldc_full == n_total) {
    // row panels are contiguous in c_full: gather straight into place
    GUDA_NCCL(g_nccl.AllGather(c_block, c_full, per_block, dt, c->comm, s));
    return GUDA_OK;
  }
  size_t need = per_block * c->nranks * sizeof(T);
  if (need > c->staging_bytes) {
    GUDA_CUDA(cudaStreamSynchronize(s));
    if (c->staging) GUDA_CUDA(cudaFree(c->staging));
    c->staging = nullptr; c->staging_bytes = 0;
    GUDA_CUDA(cudaMalloc(&c->staging, need));
    c->staging_bytes = need;
  }
  GUDA_NCCL(g_nccl.AllGather(c_block, c->staging, per_block, dt, c->comm, s));
  unpack_blocks_kernel<T, V><<<sm_count() * 8, 256, 0, s>>>(static_cast<const T*>(c->staging), c_full, ldc_full, pr, pc, m_loc, n_loc);
  return check_launch("unpack_blocks");
}

}  // namespace guda

using namespace guda;
extern "C" {

int guda_comm_unique_id(uint8_t host_id[GUDA_COMM_ID_BYTES]) {
  int rc = load_nccl();
  if (rc) return rc;
  static_assert(sizeof(ncclUniqueId) == GUDA_COMM_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId id;
  GUDA_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(host_id, &id, sizeof(id));
  return GUDA_OK;
}

int guda_comm_init_rank(guda_comm_t* comm, int nranks, int rank, const uint8_t host_id[GUDA_COMM_ID_BYTES]) {
  if (!comm || nranks < 1 || rank < 0 || rank >= nranks) return set_error(GUDA_ERR_INVALID_ARG, "comm_init_rank: bad arguments");
  int rc = load_nccl();
  if (rc) return rc;
  rc = bind_device();
  if (rc) return rc;
  ncclUniqueId id;
  memcpy(&id, host_id, sizeof(id));
  Comm* c = new Comm();
  c->nranks = nranks; c->rank = rank;
  ncclResult_t r = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
  if (r != ncclSuccess) { delete c; return set_error(GUDA_ERR_EXECUTION, "ncclCommInitRank failed: %s", g_nccl.GetErrorString(r)); }
  cudaError_t e = cudaMalloc((void**)&c->dev_scalar, 256);
  if (e == cudaSuccess) e = cudaMalloc((void**)&c->dev_records, sizeof(PeerRecord) * (nranks + 1));
  if (e == cudaSuccess) e = cudaHostAlloc((void**)&c->host_records, sizeof(PeerRecord) * (nranks + 1), cudaHostAllocDefault);
  if (e != cudaSuccess) { g_nccl.CommDestroy(c->comm); delete c; return set_error(GUDA_ERR_MEMORY, "cudaMalloc: %s", cudaGetErrorString(e)); }
  c->opened.resize(nranks);
  if (const char* m = getenv("GUDA_SHARDED_GATHER")) c->gather_mode = (strcmp(m, "nccl") == 0) ? GUDA_GATHER_NCCL : GUDA_GATHER_P2P;
  if (nranks > GUDA_MAX_SCATTER_RANKS) c->gather_mode = GUDA_GATHER_NCCL;
  *comm = reinterpret_cast<guda_comm_t>(c);
  return GUDA_OK;
}

int guda_comm_destroy(guda_comm_t comm) {
  if (!comm) return set_error(GUDA_ERR_INVALID_ARG, "null communicator");
  int rc = bind_device();
  if (rc) return rc;
  Comm* c = reinterpret_cast<Comm*>(comm);
  cudaDeviceSynchronize();
  for (auto& per_rank : c->opened)
    for (PeerMapping& pm : per_rank) cudaIpcCloseMemHandle(pm.base);
  if (c->row_comm) g_nccl.CommDestroy(c->row_comm);
  if (c->col_comm) g_nccl.CommDestroy(c->col_comm);
  g_nccl.CommDestroy(c->comm);
  if (c->dev_records) cudaFree(c->dev_records);
  if (c->host_records) cudaFreeHost(c->host_records);
  if (c->staging) cudaFree(c->staging);
  if (c->dev_scalar) cudaFree(c->dev_scalar);
  delete c;
  return GUDA_OK;
}

// beta applies to this rank's c_block (the K-chunked host form resumes the FMA chain with beta = 1); c_full receives the result
static int sgemm_sharded_impl(guda_comm_t comm, int pr, int pc, int m_total, int n_total, int k, float alpha, const float* a_panel, int lda,
                              const float* b_panel, int ldb, float beta, float* c_block, int ldc_block, float* c_full, int ldc_full,
                              guda_stream_t stream) {
  if (!comm) return set_error(GUDA_ERR_INVALID_ARG, "null communicator");
  Comm* c = reinterpret_cast<Comm*>(comm);
  if (pr < 1 || pc < 1 || pr * pc != c->nranks) return set_error(GUDA_ERR_INVALID_ARG, "process grid %dx%d does not match %d ranks", pr, pc, c->nranks);
  if (m_total < 0) GUDA_PANIC("blas: m < 0");
  if (n_total < 0) GUDA_PANIC("blas: n < 0");
  if (m_total % pr || n_total % pc) return set_error(GUDA_ERR_INVALID_ARG, "m (n) must divide evenly over the process grid");
  const int m_loc = m_total / pr, n_loc = n_total / pc;
  if (ldc_block != n_loc) return set_error(GUDA_ERR_INVALID_ARG, "c_block must be packed (ldc_block == n/pc) for the all-gather");
  int rc = bind_device();
  if (rc) return rc;
  cudaStream_t s = native(stream);
  if (c_full && ldc_full < n_total) GUDA_PANIC("blas: bad leading dimension of C");
  // ---- fused route: ONE kernel computes the block (full K locally: the same per-element k-order as on one GPU, so the
  // bits are identical) and its epilogue stores every finished tile into every rank's c_full over NVLink ----
  if (c_full && c->gather_mode == GUDA_GATHER_P2P && !c->p2p_failed && sgemm_mode() == GUDA_SGEMM_EXACT) {
    void* peers_v[GUDA_MAX_SCATTER_RANKS];
    bool ok = false;
    if ((rc = resolve_peers(c, c_full, s, peers_v, &ok))) return rc;
    if (ok) {
      float* peers[GUDA_MAX_SCATTER_RANKS];
      for (int i = 0; i < c->nranks; i++) peers[i] = static_cast<float*>(peers_v[i]);
      const int r = c->rank / pc, q = c->rank % pc;
      rc = guda_sgemm_scatter(0, 0, m_loc, n_loc, k, alpha, a_panel, lda, b_panel, ldb, beta, c_block, ldc_block, peers, c->nranks,
                              ldc_full, r * m_loc, q * n_loc, stream);
      if (rc) return rc;
      // exit barrier: the one-float all-reduce completes on this stream only after every rank's kernel has finished
      // storing into this rank's c_full
      GUDA_NCCL(g_nccl.AllReduce(c->dev_scalar, c->dev_scalar, 1, ncclFloat, ncclSum, c->comm, s));
      return GUDA_OK;
    }
    c->p2p_failed = true;                  // agreed by all ranks (resolve_peers reduces the flag): use NCCL from now on
  }
  // ---- NCCL route: local block, then all-gather ----
  rc = guda_sgemm(0, 0, m_loc, n_loc, k, alpha, a_panel, lda, b_panel, ldb, beta, c_block, ldc_block, stream);
  if (rc) return rc;
  if (!c_full) return GUDA_OK;             // caller only wants its block
  return gather_blocks_nccl<float, float4>(c, pr, pc, m_loc, n_loc, n_total, c_block, c_full, ldc_full, ncclFloat, s);
}

// DGEMM twin (row j1 of the scope table: "SGEMM/DGEMM ... @1/2/4/8 B200").  The reference's Dgemm is single-threaded
// (compute/dgemm.go:122); the 2-D partition with full K per rank keeps every element's chain, so the gathered matrix is
// bit-identical to the one-GPU product.  Fused route: the DMMA kernel's epilogue stores into every rank's c_full.
static int dgemm_sharded_impl(guda_comm_t comm, int pr, int pc, int m_total, int n_total, int k, double alpha, const double* a_panel, int lda,
                              const double* b_panel, int ldb, double beta, double* c_block, int ldc_block, double* c_full, int ldc_full,
                              guda_stream_t stream) {
  if (!comm) return set_error(GUDA_ERR_INVALID_ARG, "null communicator");
  Comm* c = reinterpret_cast<Comm*>(comm);
  if (pr < 1 || pc < 1 || pr * pc != c->nranks) return set_error(GUDA_ERR_INVALID_ARG, "process grid %dx%d does not match %d ranks", pr, pc, c->nranks);
  if (m_total < 0) GUDA_PANIC("blas: m < 0");
  if (n_total < 0) GUDA_PANIC("blas: n < 0");
  if (m_total % pr || n_total % pc) return set_error(GUDA_ERR_INVALID_ARG, "m (n) must divide evenly over the process grid");
  const int m_loc = m_total / pr, n_loc = n_total / pc;
  if (ldc_block != n_loc) return set_error(GUDA_ERR_INVALID_ARG, "c_block must be packed (ldc_block == n/pc) for the all-gather");
  int rc = bind_device();
  if (rc) return rc;
  cudaStream_t s = native(stream);
  if (c_full && ldc_full < n_total) GUDA_PANIC("blas: bad leading dimension of C");
  const bool p2p_shape = m_loc % 4 == 0 && n_loc % 4 == 0 && ldc_full % 4 == 0 && ((uintptr_t)c_full % 32) == 0 && ((uintptr_t)c_block % 32) == 0;
  if (c_full && c->gather_mode == GUDA_GATHER_P2P && !c->p2p_failed && p2p_shape) {
    void* peers_v[GUDA_MAX_SCATTER_RANKS];
    bool ok = false;
    if ((rc = resolve_peers(c, c_full, s, peers_v, &ok))) return rc;
    if (ok) {
      double* peers[GUDA_MAX_SCATTER_RANKS];
      for (int i = 0; i < c->nranks; i++) peers[i] = static_cast<double*>(peers_v[i]);
      const int r = c->rank / pc, q = c->rank % pc;
      rc = guda_dgemm_scatter(0, 0, m_loc, n_loc, k, alpha, a_panel, lda, b_panel, ldb, beta, c_block, ldc_block, peers, c->nranks,
                              ldc_full, r * m_loc, q * n_loc, stream);
      if (rc) return rc;
      GUDA_NCCL(g_nccl.AllReduce(c->dev_scalar, c->dev_scalar, 1, ncclFloat, ncclSum, c->comm, s));     // exit barrier
      return GUDA_OK;
    }
    c->p2p_failed = true;
  }
  rc = guda_dgemm(0, 0, m_loc, n_loc, k, alpha, a_panel, lda, b_panel, ldb, beta, c_block, ldc_block, stream);
  if (rc) return rc;
  if (!c_full) return GUDA_OK;
  return gather_blocks_nccl<double, double2>(c, pr, pc, m_loc, n_loc, n_total, c_block, c_full, ldc_full, ncclDouble, s);
}

int guda_sgemm_sharded(guda_comm_t comm, int pr, int pc, int m_total, int n_total, int k, float alpha, const float* a_panel, int lda,
                       const float* b_panel, int ldb, float* c_block, int ldc_block, float* c_full, int ldc_full, guda_stream_t stream) {
  return sgemm_sharded_impl(comm, pr, pc, m_total, n_total, k, alpha, a_panel, lda, b_panel, ldb, 0.0f, c_block, ldc_block, c_full, ldc_full, stream);
}

int guda_dgemm_sharded(guda_comm_t comm, int pr, int pc, int m_total, int n_total, int k, double alpha, const double* a_panel, int lda,
                       const double* b_panel, int ldb, double* c_block, int ldc_block, double* c_full, int ldc_full, guda_stream_t stream) {
  return dgemm_sharded_impl(comm, pr, pc, m_total, n_total, k, alpha, a_panel, lda, b_panel, ldb, 0.0, c_block, ldc_block, c_full, ldc_full, stream);
}

// The same step from HOST panels (each rank's A row panel and B column panel in pinned host memory) to a HOST C block: the uploads
// are pipelined under the GEMM by 8192-deep K chunks (chunk j's FMA chains resume from the stored block with beta = 1: same
// bits), the last chunk is the fused GEMM + all-gather launch, then the block returns to the host.  Blocks.
int guda_sgemm_sharded_host(guda_comm_t comm, int pr, int pc, int m_total, int n_total, int k, float alpha, const float* host_a_panel, int lda,
                            const float* host_b_panel, int ldb, float* host_c_block, int ldc_host, float* a_panel_dev, float* b_panel_dev,
                            float* c_block_dev, float* c_full, int ldc_full, guda_stream_t stream) {
  if (!comm) return set_error(GUDA_ERR_INVALID_ARG, "null communicator");
  Comm* c = reinterpret_cast<Comm*>(comm);
  if (pr < 1 || pc < 1 || pr * pc != c->nranks) return set_error(GUDA_ERR_INVALID_ARG, "process grid %dx%d does not match %d ranks", pr, pc, c->nranks);
  if (m_total < 0) GUDA_PANIC("blas: m < 0");
  if (n_total < 0) GUDA_PANIC("blas: n < 0");
  if (k < 1) return set_error(GUDA_ERR_INVALID_ARG, "sgemm_sharded_host: k must be positive");
  if (m_total % pr || n_total % pc) return set_error(GUDA_ERR_INVALID_ARG, "m (n) must divide evenly over the process grid");
  const int m_loc = m_total / pr, n_loc = n_total / pc;
  if (lda < k) GUDA_PANIC("blas: bad leading dimension of A");
  if (ldb < n_loc) GUDA_PANIC("blas: bad leading dimension of B");
  if (ldc_host < n_loc) GUDA_PANIC("blas: bad leading dimension of C");
  if (!host_a_panel || !host_b_panel || !host_c_block || !a_panel_dev || !b_panel_dev || !c_block_dev)
    return set_error(GUDA_ERR_INVALID_ARG, "sgemm_sharded_host: null pointer");
  int rc = bind_device();
  if (rc) return rc;
  cudaStream_t s = native(stream), s_in, s_in2, s_out;             // s_in carries the A chunks, s_in2 the B chunks
  if ((rc = aux_copy_streams(&s_in, &s_out, &s_in2))) return rc;
  struct EventBag {                       // destroyed on every return path (an early error return used to leak the step's events);
    std::vector<cudaEvent_t> v;            // destroying an event that queued work still references is legal: CUDA defers the release
    ~EventBag() { for (cudaEvent_t e : v) cudaEventDestroy(e); }
    void push_back(cudaEvent_t e) { v.push_back(e); }
  } evs;
  // GUDA_HOST_STEP_TRACE=1: events carry timestamps and every rank prints, per K chunk, when its operands were on the device and when
  // its GEMM finished (ms from the start of the step) -- the tool that shows whether the GEMM ever waits for the upload pipeline
  static const bool trace = [] { const char* e = getenv("GUDA_HOST_STEP_TRACE"); return e && atoi(e) != 0; }();
  auto new_event = [&](cudaEvent_t* e) -> int {
    GUDA_CUDA(cudaEventCreateWithFlags(e, trace ? cudaEventDefault : cudaEventDisableTiming));
    evs.push_back(*e);
    return GUDA_OK;
  };
  std::vector<cudaEvent_t> tr_in, tr_done;
  std::vector<double> tr_host;                                     // host clock (ms since entry) when chunk j's copies / GEMM were enqueued
  const auto t_entry = std::chrono::steady_clock::now();
  auto host_ms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_entry).count(); };
  cudaEvent_t e0;
  if ((rc = new_event(&e0))) return rc;
  GUDA_CUDA(cudaEventRecord(e0, s));
  GUDA_CUDA(cudaStreamWaitEvent(s_in, e0, 0));
  GUDA_CUDA(cudaStreamWaitEvent(s_in2, e0, 0));
  std::vector<int> kb;
  host_k_chunks(k, kb);
  const int nk = (int)kb.size() - 1;
  bool downloaded = false;
  // Every unique panel chunk crosses PCIe ONCE: the ranks of a grid row share the A row panel and the ranks of a grid column the
  // B column panel, so one of them (chosen so that no rank owns two panels when the grid allows it) uploads the chunk from its
  // host copy and ncclBroadcast fans it out over NVLink inside the row / column sub-communicator, all on the copy stream and
  // therefore under the previous chunk's GEMM.  Round 1 uploaded every panel from every rank that needed it: 96 GiB per step at
  // 8 GPUs against 32 GiB unique, and the step was bound by host memory bandwidth.
  const int gr = c->rank / pc, gc = c->rank % pc;
  if (c->nranks > 1 && (c->grid_pr != pr || c->grid_pc != pc)) {
    if (c->row_comm) { g_nccl.CommDestroy(c->row_comm); c->row_comm = nullptr; }
    if (c->col_comm) { g_nccl.CommDestroy(c->col_comm); c->col_comm = nullptr; }
    // The broadcasts run under a GEMM that owns every SM, and a rank that only receives sits in its broadcast kernel while the
    // root is still uploading: keep those kernels to a few CTAs (they need ~10 GB/s, not the link rate).  GUDA_BCAST_CTAS overrides.
    ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
    static const int bcast_ctas = [] { const char* e = getenv("GUDA_BCAST_CTAS"); return e ? atoi(e) : 4; }();
    if (bcast_ctas > 0) { cfg.minCTAs = 1; cfg.maxCTAs = bcast_ctas; }
    if (pc > 1) GUDA_NCCL(g_nccl.CommSplit(c->comm, gr, gc, &c->row_comm, &cfg));
    if (pr > 1) GUDA_NCCL(g_nccl.CommSplit(c->comm, pr + gc, gr, &c->col_comm, &cfg));
    c->grid_pr = pr; c->grid_pc = pc;
  }
  const int a_root = gr % pc, b_root = (gc + 1) % pr;            // rank inside row_comm / col_comm that uploads the shared panel
  const bool a_owner = gc == a_root, b_owner = gr == b_root;
  // device staging of A is chunk-major: chunk j is a dense m_loc x kc block at a_panel_dev + m_loc * k0 (leading dimension kc),
  // so that a chunk is ONE contiguous buffer for the broadcast; B's k-chunks are contiguous row blocks already
  for (int j = 0; j < nk && !rc; j++) {
    const int k0 = kb[j], kc = kb[j + 1] - kb[j];
    float* a_chunk = a_panel_dev + (size_t)m_loc * k0;
    float* b_chunk = b_panel_dev + (size_t)k0 * n_loc;
    if (trace) tr_host.push_back(host_ms());
    if (a_owner)
      GUDA_CUDA(cudaMemcpy2DAsync(a_chunk, (size_t)kc * 4, host_a_panel + k0, (size_t)lda * 4, (size_t)kc * 4, m_loc, cudaMemcpyHostToDevice, s_in));
    if (a_owner && b_owner) {
      // one H2D engine serves both streams in submission order: without this edge every B chunk (no dependencies among them) is
      // submitted at once and A chunk j+1, which becomes ready only after broadcast j, queues behind ALL of B (2 GPUs: 150 ms stall)
      cudaEvent_t e_a_up;
      if ((rc = new_event(&e_a_up))) break;
      GUDA_CUDA(cudaEventRecord(e_a_up, s_in));
      GUDA_CUDA(cudaStreamWaitEvent(s_in2, e_a_up, 0));
    }
    if (c->row_comm) GUDA_NCCL(g_nccl.Broadcast(a_chunk, a_chunk, (size_t)m_loc * kc, ncclFloat, a_root, c->row_comm, s_in));
    if (b_owner)
      GUDA_CUDA(cudaMemcpy2DAsync(b_chunk, (size_t)n_loc * 4, host_b_panel + (size_t)k0 * ldb, (size_t)ldb * 4, (size_t)n_loc * 4, kc,
                                  cudaMemcpyHostToDevice, s_in2));
    if (c->col_comm) GUDA_NCCL(g_nccl.Broadcast(b_chunk, b_chunk, (size_t)kc * n_loc, ncclFloat, b_root, c->col_comm, s_in2));
    cudaEvent_t e_in, e_in_b;                                       // the two operand streams never wait for each other
    if ((rc = new_event(&e_in)) || (rc = new_event(&e_in_b))) break;
    GUDA_CUDA(cudaEventRecord(e_in, s_in));
    GUDA_CUDA(cudaEventRecord(e_in_b, s_in2));
    GUDA_CUDA(cudaStreamWaitEvent(s, e_in, 0));
    GUDA_CUDA(cudaStreamWaitEvent(s, e_in_b, 0));
    if (trace) { tr_in.push_back(e_in); tr_in.push_back(e_in_b); tr_host.push_back(host_ms()); }
    const float bj = j == 0 ? 0.0f : 1.0f;
    if (j + 1 < nk) {
      rc = guda_sgemm(0, 0, m_loc, n_loc, kc, alpha, a_chunk, kc, b_chunk, n_loc, bj, c_block_dev, n_loc, stream);
      if (trace && !rc) {
        cudaEvent_t e_done;
        if ((rc = new_event(&e_done))) break;
        GUDA_CUDA(cudaEventRecord(e_done, s));
        tr_done.push_back(e_done);
      }
    } else {
      // last chunk.  Fused route: the GEMM + all-gather launch is cut into row panels so that each finished panel of this rank's
      // block streams back to the host while the next one computes; otherwise the generic step, then one download.
      bool fused = false;
      float* peers[GUDA_MAX_SCATTER_RANKS];
      if (c_full && ldc_full >= n_total && c->gather_mode == GUDA_GATHER_P2P && !c->p2p_failed && sgemm_mode() == GUDA_SGEMM_EXACT) {
        void* peers_v[GUDA_MAX_SCATTER_RANKS];
        if ((rc = resolve_peers(c, c_full, s, peers_v, &fused))) break;
        for (int i = 0; i < c->nranks; i++) peers[i] = static_cast<float*>(peers_v[i]);
        if (!fused) c->p2p_failed = true;
      }
      if (!fused) {
        rc = sgemm_sharded_impl(comm, pr, pc, m_total, n_total, kc, alpha, a_chunk, kc, b_chunk, n_loc, bj,
                                c_block_dev, n_loc, c_full, ldc_full, stream);
        break;
      }
      GUDA_CUDA(cudaStreamWaitEvent(s_out, e0, 0));
      const int r = c->rank / pc, q = c->rank % pc;
      const int panels = host_c_panels(m_loc);
      const int pm = ((m_loc + panels - 1) / panels + 127) / 128 * 128;
      for (int p0 = 0; p0 < m_loc && !rc; p0 += pm) {
        const int ph = m_loc - p0 < pm ? m_loc - p0 : pm;
        rc = guda_sgemm_scatter(0, 0, ph, n_loc, kc, alpha, a_chunk + (size_t)p0 * kc, kc, b_chunk, n_loc, bj,
                                c_block_dev + (size_t)p0 * n_loc, n_loc, peers, c->nranks, ldc_full, r * m_loc + p0, q * n_loc, stream);
        if (rc) break;
        cudaEvent_t e_panel;
        if ((rc = new_event(&e_panel))) break;
        GUDA_CUDA(cudaEventRecord(e_panel, s));
        GUDA_CUDA(cudaStreamWaitEvent(s_out, e_panel, 0));
        GUDA_CUDA(cudaMemcpy2DAsync(host_c_block + (size_t)p0 * ldc_host, (size_t)ldc_host * 4, c_block_dev + (size_t)p0 * n_loc, (size_t)n_loc * 4,
                                    (size_t)n_loc * 4, ph, cudaMemcpyDeviceToHost, s_out));
      }
      if (rc) break;
      // exit barrier (every rank's kernels have finished storing into this rank's c_full), then wait for the downloads
      GUDA_NCCL(g_nccl.AllReduce(c->dev_scalar, c->dev_scalar, 1, ncclFloat, ncclSum, c->comm, s));
      cudaEvent_t e_out;
      if ((rc = new_event(&e_out))) break;
      GUDA_CUDA(cudaEventRecord(e_out, s_out));
      GUDA_CUDA(cudaStreamWaitEvent(s, e_out, 0));
      downloaded = true;
    }
  }
  if (!rc && downloaded) {
    cudaError_t ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) rc = set_error(map_cuda_error(ce), "sgemm_sharded_host: %s", cudaGetErrorString(ce));
  } else if (!rc) {
    cudaError_t ce = cudaMemcpy2DAsync(host_c_block, (size_t)ldc_host * 4, c_block_dev, (size_t)n_loc * 4, (size_t)n_loc * 4, m_loc,
                                       cudaMemcpyDeviceToHost, s);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) rc = set_error(map_cuda_error(ce), "sgemm_sharded_host: %s", cudaGetErrorString(ce));
  }
  if (trace && !rc) {
    cudaEvent_t e_end;
    if (!new_event(&e_end) && cudaEventRecord(e_end, s) == cudaSuccess && cudaEventSynchronize(e_end) == cudaSuccess) {
      std::string line = "[guda host step] rank " + std::to_string(c->rank) + " k-bounds/in/done ms:";
      for (size_t j = 0; 2 * j + 1 < tr_in.size(); j++) {
        float t_a = 0, t_b = 0, t_done = 0;
        cudaEventElapsedTime(&t_a, e0, tr_in[2 * j]);
        cudaEventElapsedTime(&t_b, e0, tr_in[2 * j + 1]);
        if (j < tr_done.size()) cudaEventElapsedTime(&t_done, e0, tr_done[j]);
        char buf[160];
        snprintf(buf, sizeof buf, " %d:A%.1f,B%.1f/%.1f[host %.1f-%.1f]", kb[j + 1], t_a, t_b, t_done, tr_host[2 * j], tr_host[2 * j + 1]);
        line += buf;
      }
      float t_end = 0;
      cudaEventElapsedTime(&t_end, e0, e_end);
      fprintf(stderr, "%s end %.1f\n", line.c_str(), t_end);
    }
  }
  return rc;
}

int guda_comm_set_gather_mode(guda_comm_t comm, int mode) {
  if (!comm || (mode != GUDA_GATHER_P2P && mode != GUDA_GATHER_NCCL)) return set_error(GUDA_ERR_INVALID_ARG, "comm_set_gather_mode: bad arguments");
  reinterpret_cast<Comm*>(comm)->gather_mode = mode;
  return GUDA_OK;
}
int guda_comm_get_gather_mode(guda_comm_t comm, int* mode_in_effect) {
  if (!comm || !mode_in_effect) return set_error(GUDA_ERR_INVALID_ARG, "comm_get_gather_mode: bad arguments");
  Comm* c = reinterpret_cast<Comm*>(comm);
  *mode_in_effect = (c->gather_mode == GUDA_GATHER_P2P && !c->p2p_failed) ? GUDA_GATHER_P2P : GUDA_GATHER_NCCL;
  return GUDA_OK;
}

// AXPY over a vector split into contiguous chunks, one per rank (math.go:39): elementwise, so no collective at all; the
// entry point exists so that the sharded Level-1 pair has the same shape (and the same argument checks) as the DOT below.
int guda_saxpy_sharded(guda_comm_t comm, int n_local, float alpha, const float* x, float* y, guda_stream_t stream) {
  if (!comm) return set_error(GUDA_ERR_INVALID_ARG, "saxpy_sharded: null communicator");
  return guda_saxpy(n_local, alpha, x, 1, y, 1, stream);
}

// stream-ordered form: the combined scalar is written to dev_result on every rank; nothing blocks
int guda_sdot_sharded_device(guda_comm_t comm, int n_local, const float* x, const float* y, float* dev_result, guda_stream_t stream) {
  if (!comm || !dev_result) return set_error(GUDA_ERR_INVALID_ARG, "sdot_sharded: null argument");
  Comm* c = reinterpret_cast<Comm*>(comm);
  int rc = guda_sdot_device(n_local, x, 1, y, 1, dev_result, stream);
  if (rc) return rc;
  GUDA_NCCL(g_nccl.AllReduce(dev_result, dev_result, 1, ncclFloat, ncclSum, c->comm, native(stream)));
  return GUDA_OK;
}

int guda_sdot_sharded(guda_comm_t comm, int n_local, const float* x, const float* y, float* host_result, guda_stream_t stream) {
  if (!comm || !host_result) return set_error(GUDA_ERR_INVALID_ARG, "sdot_sharded: null argument");
  Comm* c = reinterpret_cast<Comm*>(comm);
  int rc = guda_sdot_device(n_local, x, 1, y, 1, c->dev_scalar, stream);
  if (rc) return rc;
  cudaStream_t s = native(stream);
  GUDA_NCCL(g_nccl.AllReduce(c->dev_scalar, c->dev_scalar, 1, ncclFloat, ncclSum, c->comm, s));
  GUDA_CUDA(cudaMemcpyAsync(host_result, c->dev_scalar, sizeof(float), cudaMemcpyDeviceToHost, s));
  GUDA_CUDA(cudaStreamSynchronize(s));
  return GUDA_OK;
}

}  // extern "C"
