// host_gemm.cu -- SGEMM on HOST-resident operands: the end-to-end form of math.go:98 GEMM for callers whose matrices live in
// host memory (what the reference's DevicePtr always was: memory.go:246-300 handed out host slices).
//
// The copies are not bracketed around the GEMM, they are pipelined under it.  An fp32 C element is ONE sequential FMA chain
// (compute/sgemm.go:252-260), so the chain can be run in K chunks -- C stored after chunk j, resumed with beta = 1 -- without
// changing a bit; that is the same property the device path already uses for its 8192-deep K-blocked launches.  So:
//   copy-in stream : A[:, chunk j] and B[chunk j, :] for j = 0, 1, ...            (pinned host -> device, 2-D DMA)
//   compute stream : chunk j's GEMM as soon as its operands have landed; chunk j+1 uploads meanwhile
//   copy-out stream: the LAST chunk runs as row panels; each finished panel of C goes back to the host while the next computes
// Exposed transfer time is one chunk in and one panel out instead of the whole of A, B and C.
// Results are bit-identical to guda_sgemm on device-resident operands (tests/test_gpu_host_gemm.py).
#include "common.cuh"
#include <mutex>
#include <vector>

namespace guda {

constexpr int HG_KC = 8192;          // = the device path's K block (gemm.cu KChunk<float>): each chunk is a single launch

struct HgStreams { cudaStream_t in = nullptr, in2 = nullptr, out = nullptr; };
static int hg_streams(HgStreams** o) {
  static HgStreams st;
  static std::mutex mu;                    // first use may come from several host threads at once
  std::lock_guard<std::mutex> lk(mu);
  if (!st.out) {                           // `out` is created last: a failed attempt is simply repeated by the next call
    // Highest priority: what runs on these streams besides copies is the NVLink broadcast of the sharded host step (an NCCL kernel).
    // The GEMM launches fill every SM (4 CTAs x 126 registers x 128 threads), and at equal priority a waiting kernel only gets CTA
    // slots when the running grid has none left to issue -- one broadcast per GEMM launch boundary, which starved the upload
    // pipeline at 8 GPUs.  At higher priority its few CTAs take the first slots a retiring GEMM CTA frees (every ~0.1 ms).
    int least = 0, greatest = 0;
    GUDA_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    if (!st.in) GUDA_CUDA(cudaStreamCreateWithPriority(&st.in, cudaStreamNonBlocking, greatest));
    if (!st.in2) GUDA_CUDA(cudaStreamCreateWithPriority(&st.in2, cudaStreamNonBlocking, greatest));
    GUDA_CUDA(cudaStreamCreateWithPriority(&st.out, cudaStreamNonBlocking, greatest));
  }
  *o = &st;
  return GUDA_OK;
}

// K-chunk schedule: a short first chunk (its upload is the only one nothing can hide), then doubling depths up to HG_KC, then
// HG_KC-deep chunks: 0, 512, 1024, 2048, 4096, 8192, 16384, ...  Chunk j+1 uploads under chunk j's GEMM, so its depth may exceed
// chunk j's only by the ratio of GEMM time to upload time per unit of k -- about 13 on one GPU, but only 2-3 at 8 GPUs, where a
// rank's GEMM share is 1/8 while the rank that uploads a shared A row panel still moves half of A (round 2's 1024 -> 8192 jump left
// the 8-GPU GEMM waiting ~40 ms for its second chunk).  Any split keeps the bits: every chunk resumes the same FMA chains.
void host_k_chunks(int k, std::vector<int>& bounds) {
  bounds.clear();
  bounds.push_back(0);
  if (k > HG_KC)
    for (int b = HG_KC / 16; b < HG_KC && b < k; b *= 2) bounds.push_back(b);
  for (int b = HG_KC; b < k; b += HG_KC) bounds.push_back(b);
  bounds.push_back(k);
  // The LAST chunk is the only one whose results are final, so the download of C can only hide under it: make it two blocks deep
  // (8 GPUs: each rank's 2 GiB block leaves at ~12 GB/s when all eight write host memory at once, 175 ms against 130 ms of GEMM per
  // 8192-deep chunk -- the trace showed the step ending 40-75 ms after its last kernel).
  if (bounds.size() >= 5 && bounds[bounds.size() - 2] >= 2 * HG_KC) bounds.erase(bounds.end() - 2);
}
int host_c_panels(int m) { return m >= 16384 ? 16 : (m >= 2048 ? 8 : 1); }   // row panels of the last chunk: the last panel's download is the exposed tail

int aux_copy_streams(cudaStream_t* in, cudaStream_t* out, cudaStream_t* in2) {
  HgStreams* st;
  int rc = hg_streams(&st);
  if (rc) return rc;
  *in = st->in;
  *out = st->out;
  if (in2) *in2 = st->in2;
  return GUDA_OK;
}

struct HgBuffers {
  void *a = nullptr, *b = nullptr, *c = nullptr;
  std::vector<cudaEvent_t> ev;
  std::vector<cudaStream_t> used;          // streams that may still hold copies / kernels on the staging buffers
  bool done = false;                       // set once the caller's stream has been synchronised
  ~HgBuffers() {
    if (!done) for (cudaStream_t q : used) cudaStreamSynchronize(q);     // error return: drain before the buffers and events go
    for (cudaEvent_t e : ev) cudaEventDestroy(e);
    if (a) guda_free(a);
    if (b) guda_free(b);
    if (c) guda_free(c);
  }
  int event(cudaEvent_t* e) {
    GUDA_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    ev.push_back(*e);
    return GUDA_OK;
  }
};

}  // namespace guda

using namespace guda;
extern "C" int guda_sgemm_host(int ta, int tb, int m, int n, int k, float alpha, const float* host_a, int lda, const float* host_b, int ldb,
                               float beta, float* host_c, int ldc, guda_stream_t stream) {
  // dimension / leading-dimension panics in the reference's order and with its strings (compute/sgemm.go:30-71)
  if (m < 0) GUDA_PANIC("blas: m < 0");
  if (n < 0) GUDA_PANIC("blas: n < 0");
  if (k < 0) GUDA_PANIC("blas: k < 0");
  const int rows_a = ta ? k : m, cols_a = ta ? m : k, rows_b = tb ? n : k, cols_b = tb ? k : n;
  if (lda < (cols_a > 1 ? cols_a : 1)) GUDA_PANIC("blas: bad leading dimension of A");
  if (ldb < (cols_b > 1 ? cols_b : 1)) GUDA_PANIC("blas: bad leading dimension of B");
  if (ldc < (n > 1 ? n : 1)) GUDA_PANIC("blas: bad leading dimension of C");
  if (m == 0 || n == 0) return GUDA_OK;
  int rc = bind_device();
  if (rc) return rc;
  if (!host_c || ((!host_a || !host_b) && k > 0)) return set_error(GUDA_ERR_INVALID_ARG, "GEMM: null pointer");
  cudaStream_t s = native(stream);
  HgStreams* st;
  if ((rc = hg_streams(&st))) return rc;
  HgBuffers buf;
  buf.used = {st->in, st->out, s};
  // compute/sgemm.go:80-96 takes its 4^3 / 8^3 / 16^3 kernels (a different summation order) only for CONTIGUOUS operands; the staging
  // copies are dense, so when the caller's were not, a padded device ldc keeps guda_sgemm on the general path the reference takes
  const bool small_strided = !ta && !tb && m == n && n == k && (m == 4 || m == 8 || m == 16) && !(lda == k && ldb == n && ldc == n);
  const int ldd = small_strided ? n + 4 : n;                  // device leading dimension of C
  const int lda_d = cols_a > 1 ? cols_a : 1, ldb_d = cols_b > 1 ? cols_b : 1;
  const size_t fa = (size_t)rows_a * cols_a, fb = (size_t)rows_b * cols_b, fc = (size_t)m * ldd;
  if (k > 0) {
    if ((rc = guda_malloc(&buf.a, fa * sizeof(float)))) return rc;
    if ((rc = guda_malloc(&buf.b, fb * sizeof(float)))) return rc;
  }
  if ((rc = guda_malloc(&buf.c, fc * sizeof(float)))) return rc;
  float* da = static_cast<float*>(buf.a);
  float* db = static_cast<float*>(buf.b);
  float* dc = static_cast<float*>(buf.c);

  cudaEvent_t e_start;
  if ((rc = buf.event(&e_start))) return rc;
  GUDA_CUDA(cudaEventRecord(e_start, s));                     // copies must not overtake work already queued on the caller's stream
  GUDA_CUDA(cudaStreamWaitEvent(st->in, e_start, 0));
  GUDA_CUDA(cudaStreamWaitEvent(st->out, e_start, 0));
  if (beta != 0.0f)
    GUDA_CUDA(cudaMemcpy2DAsync(dc, (size_t)ldd * 4, host_c, (size_t)ldc * 4, (size_t)n * 4, m, cudaMemcpyHostToDevice, st->in));

  // K chunks (NN only: there a chunk is a column block of A and a row block of B); other forms upload whole operands
  const bool chunked = !ta && !tb && k > HG_KC;
  std::vector<int> kb;
  if (chunked) host_k_chunks(k, kb);
  const int nk = chunked ? (int)kb.size() - 1 : 1;
  for (int j = 0; j < nk; j++) {
    const int k0 = chunked ? kb[j] : 0;
    const int kc = chunked ? kb[j + 1] - kb[j] : k;
    if (k > 0) {
      if (chunked) {
        GUDA_CUDA(cudaMemcpy2DAsync(da + k0, (size_t)k * 4, host_a + k0, (size_t)lda * 4, (size_t)kc * 4, m, cudaMemcpyHostToDevice, st->in));
        GUDA_CUDA(cudaMemcpy2DAsync(db + (size_t)k0 * n, (size_t)n * 4, host_b + (size_t)k0 * ldb, (size_t)ldb * 4, (size_t)n * 4, kc,
                                    cudaMemcpyHostToDevice, st->in));
      } else {
        GUDA_CUDA(cudaMemcpy2DAsync(da, (size_t)cols_a * 4, host_a, (size_t)lda * 4, (size_t)cols_a * 4, rows_a, cudaMemcpyHostToDevice, st->in));
        GUDA_CUDA(cudaMemcpy2DAsync(db, (size_t)cols_b * 4, host_b, (size_t)ldb * 4, (size_t)cols_b * 4, rows_b, cudaMemcpyHostToDevice, st->in));
      }
    }
    cudaEvent_t e_in;
    if ((rc = buf.event(&e_in))) return rc;
    GUDA_CUDA(cudaEventRecord(e_in, st->in));
    GUDA_CUDA(cudaStreamWaitEvent(s, e_in, 0));
    const float bj = j == 0 ? beta : 1.0f;
    if (j + 1 < nk) {
      if ((rc = guda_sgemm(0, 0, m, n, kc, alpha, da + k0, k, db + (size_t)k0 * n, n, bj, dc, ldd, stream))) return rc;
      continue;
    }
    // last (or only) chunk: row panels, each copied out as soon as it is finished
    int panels = host_c_panels(m);
    int pm = ((m + panels - 1) / panels + 127) / 128 * 128;
    for (int r0 = 0; r0 < m; r0 += pm) {
      const int pr = m - r0 < pm ? m - r0 : pm;
      const float* ap = ta ? da + r0 : da + (size_t)r0 * cols_a + k0;
      if ((rc = guda_sgemm(ta, tb, pr, n, kc, alpha, ap, lda_d, chunked ? db + (size_t)k0 * n : db, ldb_d, bj, dc + (size_t)r0 * ldd, ldd, stream)))
        return rc;
      cudaEvent_t e_panel;
      if ((rc = buf.event(&e_panel))) return rc;
      GUDA_CUDA(cudaEventRecord(e_panel, s));
      GUDA_CUDA(cudaStreamWaitEvent(st->out, e_panel, 0));
      GUDA_CUDA(cudaMemcpy2DAsync(host_c + (size_t)r0 * ldc, (size_t)ldc * 4, dc + (size_t)r0 * ldd, (size_t)ldd * 4, (size_t)n * 4, pr,
                                  cudaMemcpyDeviceToHost, st->out));
    }
  }
  cudaEvent_t e_done;
  if ((rc = buf.event(&e_done))) return rc;
  GUDA_CUDA(cudaEventRecord(e_done, st->out));
  GUDA_CUDA(cudaStreamWaitEvent(s, e_done, 0));               // the caller's stream completes only when C is on the host
  GUDA_CUDA(cudaStreamSynchronize(s));                        // blocking, like the reference's GEMM; staging returns to the pool
  buf.done = true;
  return GUDA_OK;
}

// Test support (host logic only, no device call): the K-chunk schedule and the row-panel count of the host-operand GEMMs.
// bounds[0..*count-1] receives 0 = b0 < b1 < ... = k; returns GUDA_ERR_INVALID_ARG if `cap` is too small.
extern "C" int guda_host_gemm_schedule(int k, int m, int* bounds, int cap, int* count, int* panels) {
  if (k < 1 || !bounds || !count) return guda::set_error(GUDA_ERR_INVALID_ARG, "host_gemm_schedule: bad arguments");
  std::vector<int> kb;
  guda::host_k_chunks(k, kb);
  if ((int)kb.size() > cap) return guda::set_error(GUDA_ERR_INVALID_ARG, "host_gemm_schedule: %d bounds do not fit %d slots", (int)kb.size(), cap);
  for (size_t i = 0; i < kb.size(); i++) bounds[i] = kb[i];
  *count = (int)kb.size();
  if (panels) *panels = guda::host_c_panels(m);
  return GUDA_OK;
}
