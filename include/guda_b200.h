/*
 * guda_b200.h -- C ABI of the B200-native replacement for GUDA's BLAS hot path.
 *
 * This is the drop-in boundary: exactly what a cgo shim behind GUDA's Go API
 * (package guda, root directory of LynnColeArt/guda) binds.  Each entry point
 * cites the reference interface it replaces (file:line in the reference).
 * All pointers are DEVICE pointers unless the parameter name starts with host_.
 * Every function returns a guda_status_t (0 = OK); guda_last_error() returns a
 * thread-local message.  For BLAS argument violations the message is the exact
 * Go panic string of compute/errors.go:9-35 (e.g. "blas: m < 0"), so the Go shim
 * can re-panic with the reference's text.
 *
 * Matrices are ROW-MAJOR with ld = row stride, as in gonum/GUDA
 * (compute/sgemm.go:51-71).  There is no CPU fallback: without a CUDA device
 * every compute entry point fails with GUDA_ERR_DEVICE.
 */
#ifndef GUDA_B200_H
#define GUDA_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* errors.go:11-24 ErrorType (+1 so that 0 is success) */
typedef enum {
  GUDA_OK = 0,
  GUDA_ERR_MEMORY = 1,          /* ErrTypeMemory */
  GUDA_ERR_INVALID_ARG = 2,     /* ErrTypeInvalidArg */
  GUDA_ERR_EXECUTION = 3,       /* ErrTypeExecution */
  GUDA_ERR_NUMERICAL = 4,       /* ErrTypeNumerical */
  GUDA_ERR_DEVICE = 5,          /* ErrTypeDevice */
  GUDA_ERR_NOT_IMPLEMENTED = 6  /* ErrTypeNotImplemented; also ErrNotSupported (fused.go:11-17) */
} guda_status_t;

/* memory.go:13-21 MemcpyKind (iota order) */
typedef enum {
  GUDA_MEMCPY_H2H = 0, GUDA_MEMCPY_H2D = 1, GUDA_MEMCPY_D2H = 2, GUDA_MEMCPY_D2D = 3, GUDA_MEMCPY_DEFAULT = 4
} guda_memcpy_kind_t;

typedef struct guda_stream_s* guda_stream_t;   /* NULL = the default stream (guda.go:124-127) */
typedef struct guda_event_s*  guda_event_t;

typedef struct {                /* guda.go:34-40 Device, extended with what a GPU has */
  int    id;
  char   name[128];
  size_t total_mem;
  int    sm_count;              /* Device.NumCores */
  int    max_threads;           /* Device.MaxThreads: sm_count * max threads per SM */
  int    cc_major, cc_minor;
  int    l2_bytes;
} guda_device_props_t;

const char* guda_last_error(void);
const char* guda_version(void);

/* ---- device / context: guda.go:109-128 init(), :209-237 GetDevice/SetDevice/GetDeviceCount/GetDeviceProperties ---- */
int guda_init(int device);
int guda_device_count(int* count);
int guda_set_device(int device);
int guda_get_device(int* device);
int guda_device_properties(int device, guda_device_props_t* props);
int guda_device_synchronize(void);                        /* guda.go:278-283 Context.Synchronize */
int guda_mem_get_info(size_t* free_bytes, size_t* total_bytes);   /* docs/manual/07-runtime-api.md MemGetInfo */

/* ---- memory: memory.go:57-66 Malloc/Free, :150-234 MemoryPool, :81-145 Memcpy ----
 * Pool semantics kept: size rounded up (here to 256 B, the cudaMalloc granule; the reference
 * rounds to 64 B), first-fit reuse of freed blocks, unknown pointer -> MEMORY error,
 * second free -> MEMORY error "double free detected", stats: allocated falls on free, peak never. */
int guda_malloc(void** dptr, size_t size);
int guda_free(void* dptr);
int guda_pool_stats(int64_t* allocated, int64_t* peak);   /* memory.go:230-234 GetStats */
int guda_pool_trim(void);                                 /* new: return cached blocks to the driver */
int guda_malloc_host(void** host_ptr, size_t size);       /* pinned staging for H2D/D2H */
int guda_free_host(void* host_ptr);
int guda_memcpy(void* dst, const void* src, size_t size, int kind, guda_stream_t stream);       /* blocking, like the reference */
int guda_memcpy_async(void* dst, const void* src, size_t size, int kind, guda_stream_t stream); /* manual: MemcpyAsync */
int guda_memset(void* dptr, int value, size_t size, guda_stream_t stream);                      /* manual: Memset */

/* ---- streams and events: guda.go:57-62,242-305 Stream; docs/manual/07-runtime-api.md:122,179-192 ---- */
int guda_stream_create(guda_stream_t* stream);
int guda_stream_destroy(guda_stream_t stream);
int guda_stream_synchronize(guda_stream_t stream);        /* Stream.Synchronize guda.go:297-299 */
int guda_stream_query(guda_stream_t stream, int* done);
int guda_stream_wait_event(guda_stream_t stream, guda_event_t event);
int guda_stream_native(guda_stream_t stream, void** cuda_stream);   /* the cudaStream_t behind a handle (for timing/interop) */
int guda_event_create(guda_event_t* event);
int guda_event_destroy(guda_event_t event);
int guda_event_record(guda_event_t event, guda_stream_t stream);
int guda_event_synchronize(guda_event_t event);
int guda_event_query(guda_event_t event, int* done);
int guda_event_elapsed_ms(guda_event_t start, guda_event_t stop, float* ms);

/* ---- Level 1, fp32: math.go:39,64,170 AXPY/DOT/Scale -> compute/level1float32.go:225,619, level1float32_sdot.go:18 ----
 * Stream-ordered and asynchronous, except *_dot with a host result, which blocks until the scalar is on the host. */
int guda_saxpy(int n, float alpha, const float* x, int incx, float* y, int incy, guda_stream_t stream);
int guda_sscal(int n, float alpha, float* x, int incx, guda_stream_t stream);
int guda_sdot(int n, const float* x, int incx, const float* y, int incy, float* host_result, guda_stream_t stream);
int guda_sdot_device(int n, const float* x, int incx, const float* y, int incy, float* dev_result, guda_stream_t stream);
/* guda_sdot_device with NO allocation anywhere on the call path.  It has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same kernel and arithmetic as guda_sdot_device; the kernel's scratch (one fp32 partial per
 * CTA + the ticket counter, 4352 bytes) is a file-scope __device__ array that every GPU loading the library holds until exit,
 * instead of the per-stream buffers the ordinary entry points cudaMalloc on first use.  Limits and rules:
 *   - n <= GUDA_SDOT_STATIC_MAX_N, otherwise GUDA_ERR_INVALID_ARG ("... exceeds the maximum ...") and nothing is launched;
 *   - the library must already be initialised (guda_init allocates; this call will not): GUDA_ERR_INVALID_ARG otherwise;
 *   - x, y and dev_result are the caller's device buffers; the result is stream-ordered on `stream`, never copied to the host;
 *   - all calls on a device share the one scratch array (the ticket is cleared before every launch): calls that overlap on the
 *     GPU -- different streams -- race on it and return garbage.  Serialise them on one stream. */
#define GUDA_SDOT_STATIC_MAX_N (1 << 28)   /* elements per operand = the reference's own cap on an fp32 DevicePtr view (memory.go:250) */
int guda_sdot_device_static(int n, const float* x, int incx, const float* y, int incy, float* dev_result, guda_stream_t stream);
/* ---- Level 1, fp64: compute/level1float64.go:211,595, level1float64_ddot.go:14 ---- */
int guda_daxpy(int n, double alpha, const double* x, int incx, double* y, int incy, guda_stream_t stream);
int guda_dscal(int n, double alpha, double* x, int incx, guda_stream_t stream);
int guda_ddot(int n, const double* x, int incx, const double* y, int incy, double* host_result, guda_stream_t stream);
int guda_ddot_device(int n, const double* x, int incx, const double* y, int incy, double* dev_result, guda_stream_t stream);

/* ---- Level 3: math.go:98 GEMM -> compute/sgemm.go:29 Sgemm; compute/dgemm.go:25 Dgemm ----
 * C = alpha*op(A)*op(B) + beta*C, trans flags 0/1.  SGEMM math modes:
 *   GUDA_SGEMM_EXACT  exact-FP32 SIMT; NN/TN/TT forms reproduce the reference's per-element
 *                     sequential-k FMA (or unfused, TT) chain bit-for-bit
 *   GUDA_SGEMM_3XTF32 tcgen05 tensor cores, 3xTF32 split with fp32 TMEM accumulation (stated tolerance) */
enum { GUDA_SGEMM_EXACT = 0, GUDA_SGEMM_3XTF32 = 1 };
int guda_set_sgemm_mode(int mode);
/* exact-route kernel choice (A/B and test knob; results are bit-identical either way): cfg -1 automatic, 0 register-staged
 * kernel only, 1.. a forced configuration of the TMA-fed pipeline (sgemm_pipe.cu); lag < 0 keeps the current refill distance */
int guda_set_sgemm_pipe(int cfg, int lag);
int guda_get_sgemm_mode(int* mode);
int guda_sgemm(int trans_a, int trans_b, int m, int n, int k, float alpha, const float* a, int lda,
               const float* b, int ldb, float beta, float* c, int ldc, guda_stream_t stream);
int guda_dgemm(int trans_a, int trans_b, int m, int n, int k, double alpha, const double* a, int lda,
               const double* b, int ldb, double beta, double* c, int ldc, guda_stream_t stream);
/* DGEMM route: TENSOR = FP64 tensor pipe (DMMA m8n8k4, dgemm_dmma.cu) for the NN / TN forms with 16-byte aligned operands
 * (default; same bits as the SIMT route -- see the header of dgemm_dmma.cu); SIMT = DFMA kernel for every form */
enum { GUDA_DGEMM_TENSOR = 0, GUDA_DGEMM_SIMT = 1 };
int guda_set_dgemm_mode(int mode);
int guda_get_dgemm_mode(int* mode);
/* SGEMM (exact route) whose epilogue stores the finished m x n block to c (may be NULL when beta == 0) AND at (row0, col0)
 * of each of the ndest <= 8 row-major matrices host_dests[i] (leading dimension ld_dest).  The destinations may be peer-GPU
 * memory mapped through CUDA IPC: this is the kernel behind the fused GEMM + all-gather of guda_sgemm_sharded. */
int guda_sgemm_scatter(int trans_a, int trans_b, int m, int n, int k, float alpha, const float* a, int lda, const float* b, int ldb,
                       float beta, float* c, int ldc, float* const* host_dests, int ndest, int ld_dest, int row0, int col0,
                       guda_stream_t stream);
/* DGEMM twin (NN form, m and n multiples of 4, 32-byte aligned C rows and destinations): the FP64 tensor-pipe kernel's epilogue
 * performs the same peer stores; the kernel behind guda_dgemm_sharded. */
int guda_dgemm_scatter(int trans_a, int trans_b, int m, int n, int k, double alpha, const double* a, int lda, const double* b, int ldb,
                       double beta, double* c, int ldc, double* const* host_dests, int ndest, int ld_dest, int row0, int col0,
                       guda_stream_t stream);
/* bfloat16.go:104 GEMMBFloat16 (trans flags ignored by the reference -> must be 0 here), float16.go:201 GEMMFloat16.
 * 16-bit storage, fp32 accumulation; a/b/c are uint16 bit patterns. */
int guda_gemm_bf16(int trans_a, int trans_b, int m, int n, int k, float alpha, const uint16_t* a, int lda,
                   const uint16_t* b, int ldb, float beta, uint16_t* c, int ldc, guda_stream_t stream);
int guda_gemm_f16(int trans_a, int trans_b, int m, int n, int k, float alpha, const uint16_t* a, int lda,
                  const uint16_t* b, int ldb, float beta, uint16_t* c, int ldc, guda_stream_t stream);

/* 16-bit GEMM route: TENSOR = tcgen05 tensor cores when the shape is eligible (default); EXACT = upconvert +
 * exact-order fp32 SIMT + truncate, reproducing the reference's bits */
enum { GUDA_GEMM16_TENSOR = 0, GUDA_GEMM16_EXACT = 1 };
int guda_set_gemm16_mode(int mode);
int guda_get_gemm16_mode(int* mode);

/* ---- fused GEMM + bias + activation: fused.go:313 FusedGEMMBiasReLU, fused_gelu.go:65 FusedGEMMBiasGELU ----
 * C = act(alpha*A*B + bias[j]) with the epilogue applied in registers before the single store of C.
 * trans_a/trans_b != 0 or beta != 0 -> GUDA_ERR_NOT_IMPLEMENTED (ErrNotSupported, fused.go:324-329). */
int guda_sgemm_bias_relu(int trans_a, int trans_b, int m, int n, int k, float alpha, const float* a, int lda,
                         const float* b, int ldb, float beta, float* c, int ldc, const float* bias, guda_stream_t stream);
int guda_sgemm_bias_gelu(int trans_a, int trans_b, int m, int n, int k, float alpha, const float* a, int lda,
                         const float* b, int ldb, float beta, float* c, int ldc, const float* bias, guda_stream_t stream);
int guda_gemm_bf16_bias_relu(int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb,
                             uint16_t* c, int ldc, const float* bias, guda_stream_t stream);
int guda_gemm_bf16_bias_gelu(int m, int n, int k, float alpha, const uint16_t* a, int lda, const uint16_t* b, int ldb,
                             uint16_t* c, int ldc, const float* bias, guda_stream_t stream);
/* standalone epilogue: fused_gelu.go:12 AddBiasGELU, fused.go:220 AddBiasReLU: x[i] = act(x[i] + bias[i % bias_len]) */
int guda_add_bias_gelu(float* x, const float* bias, int n, int bias_len, guda_stream_t stream);
int guda_add_bias_relu(float* x, const float* bias, int n, int bias_len, guda_stream_t stream);

/* ---- SGEMM on HOST-resident operands: math.go:98 GEMM end to end for callers whose matrices are host memory (the reference's
 * DevicePtr was a host slice, memory.go:246-300).  Same arguments and panics as guda_sgemm, but a, b, c are HOST pointers (pinned
 * memory from guda_malloc_host makes the copies asynchronous).  Uploads are pipelined under the GEMM by K chunks (a C element is
 * one sequential FMA chain, so resuming it chunk by chunk keeps every bit) and C returns row panel by row panel while the next
 * panel computes.  Blocks until C is on the host.  Bit-identical to guda_sgemm on device-resident copies of the same operands. */
int guda_sgemm_host(int trans_a, int trans_b, int m, int n, int k, float alpha, const float* host_a, int lda, const float* host_b, int ldb,
                    float beta, float* host_c, int ldc, guda_stream_t stream);
/* Test support, host logic only (no device call): the K-chunk schedule guda_sgemm_host / guda_sgemm_sharded_host walk for depth k
 * (bounds[0] = 0 < ... < bounds[*count-1] = k: a short first chunk, doubling depths, 8192-deep blocks, a two-block last chunk -- the
 * only one the download of C can hide under) and the number of row panels the last chunk of an m-row block is cut into. */
int guda_host_gemm_schedule(int k, int m, int* bounds, int cap, int* count, int* panels);

/* ---- Level 2: compute/level2float32.go:244-308 Sgemv, :24-65 Sger (row-major; lda = row stride);
 * kernels compute/asm/f32/gemv.go:12-91 (GemvN / GemvT) and compute/asm/f32/ge_amd64.s (Ger).  trans: 0 = NoTrans,
 * non-zero = Trans / ConjTrans.  Dimension, ld and zero-increment panics come back as GUDA_ERR_INVALID_ARG with the
 * reference's strings; slice-length panics are the host shim's (it knows len()).  Results are bit-identical to the reference's
 * AVX2 route for every increment sign.  One reference quirk lives in the host shim because it needs len(y): GemvT with
 * beta == 0 and incY == 1 clears y[n:len(y)] too (gemv.go:61-64). */
int guda_sgemv(int trans, int m, int n, float alpha, const float* a, int lda, const float* x, int incx, float beta, float* y, int incy,
               guda_stream_t stream);
int guda_sger(int m, int n, float alpha, const float* x, int incx, const float* y, int incy, float* a, int lda, guda_stream_t stream);
/* fp64 twins: compute/level2float64.go:232-300 Dgemv, :24-65 Dger; kernels compute/asm/f64/gemvN_avx2_amd64.s, gemvT_amd64.s,
 * ger_amd64.s.  Bit-identical to the reference's AVX2+FMA route (4 FMA lanes per row for NoTrans, unfused chains otherwise). */
int guda_dgemv(int trans, int m, int n, double alpha, const double* a, int lda, const double* x, int incx, double beta, double* y, int incy,
               guda_stream_t stream);
int guda_dger(int m, int n, double alpha, const double* x, int incx, const double* y, int incy, double* a, int lda, guda_stream_t stream);

/* ---- Reductions: reduce.go:13-56 DevicePtr.Sum / SumSquares / Max / Min / ArgMax / ArgMin over x[0..n)
 * (kernels compute/asm/f32/sum_amd64.s, reduce_avx2_amd64.s) and Isamax (compute/level1float32.go:89-130).  Results go to HOST
 * memory and the calls block, like the reference's scalar returns.  n == 0: 0, 0, -Inf, +Inf, -1, -1 (reduce.go:18-44).
 * ArgMax / ArgMin / Isamax / Max / Min are bit-identical to the reference, NaN and signed-zero behaviour included; Sum and
 * SumSquares use a pairwise tree and are matched under ReduceArchTolerance (tolerance_arch.go:99-106) like DOT. */
/* Dsdot / Sdsdot: compute/level1float32_dsdot.go:18-54, level1float32_sdsdot.go:18-54 -- fp32 vectors, fp64
 * accumulation (exact products); Sdsdot = alpha + float32(Dsdot), 0 for n == 0.  Host result, blocking. */
int guda_dsdot(int n, const float* x, int incx, const float* y, int incy, double* host_result, guda_stream_t stream);
int guda_sdsdot(int n, float alpha, const float* x, int incx, const float* y, int incy, float* host_result, guda_stream_t stream);
int guda_ssum(long long n, const float* x, float* host_result, guda_stream_t stream);
int guda_ssumsq(long long n, const float* x, float* host_result, guda_stream_t stream);
int guda_smax(long long n, const float* x, float* host_result, guda_stream_t stream);
int guda_smin(long long n, const float* x, float* host_result, guda_stream_t stream);
int guda_sargmax(long long n, const float* x, long long* host_result, guda_stream_t stream);
int guda_sargmin(long long n, const float* x, long long* host_result, guda_stream_t stream);
int guda_isamax(int n, const float* x, int incx, long long* host_result, guda_stream_t stream);
/* reduce.go:180-187 Softmax (in place; compute/asm/f32/reduce.go:130-186: MaxAVX2, fastExp, sum, ScalUnitary) and fused.go:270-310
 * LayerNorm (in place; gamma / beta applied only when both are non-NULL).  The exp, normalise and scale steps follow the reference
 * operation by operation; the sums are pairwise (ReduceArchTolerance).  Both block. */
int guda_softmax(float* x, long long n, guda_stream_t stream);
int guda_layer_norm(float* x, const float* gamma, const float* beta, long long n, float eps, guda_stream_t stream);

/* reduce.go:105-178 ReduceSum / ReduceMax over a row-major m x n device matrix: axis 0 -> out[n] (chain over rows), axis 1 ->
 * out[m] (chain over columns).  Each output is one sequential chain in the reference's order (bit-identical); device result,
 * stream-ordered.  Any other axis: GUDA_ERR_INVALID_ARG "invalid axis" (ErrInvalidAxis). */
int guda_reduce_sum_axis(const float* x, int m, int n, int axis, float* out, guda_stream_t stream);
int guda_reduce_max_axis(const float* x, int m, int n, int axis, float* out, guda_stream_t stream);

/* reduce.go:247-291 TopK: the k largest values of x[0..n) in descending order with their indices, produced by the reference's
 * k rounds of selection sort (ties and NaNs come out in its order: bit-exact indices).  k is clamped to n; k <= 0 is
 * GUDA_ERR_INVALID_ARG "invalid parameter" (ErrInvalidParameter).  HOST results (min(k, n) entries), blocking;
 * cost k streaming passes over a working copy. */
int guda_topk(const float* x, long long n, int k, float* host_values, long long* host_indices, guda_stream_t stream);

/* ---- Elementwise device ops: math.go:133-169 Add / Multiply (c = a op b), :230-272 ReLU / Sigmoid (in place; Sigmoid is
 * activations.go:9-27 SigmoidFloat32 operation by operation).  Stream-ordered, bit-identical to the reference. */
int guda_add(const float* a, const float* b, float* c, long long n, guda_stream_t stream);
int guda_multiply(const float* a, const float* b, float* c, long long n, guda_stream_t stream);
int guda_relu(float* x, long long n, guda_stream_t stream);
int guda_sigmoid(float* x, long long n, guda_stream_t stream);
/* Test support: geluFloat32 and SigmoidFloat32 are functions of one float, and the kernels replace their IEEE divisions by
 * shorter correctly-rounding sequences (csrc/gemm_simt.cuh div_const_rn / div_normal_rn).  This call evaluates the sequences and
 * the plain div.rn restatement of fused_gelu.go:44-56 / activations.go:9-27 on ALL 2^32 inputs on the device; counts[0] / [1] =
 * inputs whose results differ in any bit (GELU / Sigmoid), first_bad[0] / [1] = the smallest such input.  Blocking. */
int guda_selftest_activations(unsigned long long counts[2], unsigned int first_bad[2]);
/* fused.go:18-216 FusedKernel: x[idx] is read once, the ops are applied to the value in order, and it is written once.
 * Types follow fused.go:34-42 (FusedCustom is available for the one function the reference passes to it, geluFloat32,
 * fused.go:261-267).  A binary op (Add: val += alpha * other, Mul: val *= other) consumes the next `others` pointer and is
 * skipped when none is left (fused.go:148,170).  broadcast_size > 0 selects other[(idx / broadcast_size) % broadcast_dim]
 * (fused.go:152-161 with broadcast_size = prod(shape[dim:]), broadcast_dim = shape[dim]); 0 selects other[idx].
 * LinearReLU (fused.go:252) = {MulScalar alpha, AddScalar beta, ReLU}; GELU (fused.go:261) = {Gelu}. */
enum { GUDA_FUSED_ADD = 0, GUDA_FUSED_MUL = 1, GUDA_FUSED_ADD_SCALAR = 2, GUDA_FUSED_MUL_SCALAR = 3, GUDA_FUSED_RELU = 4,
       GUDA_FUSED_SIGMOID = 5, GUDA_FUSED_GELU = 6 };
#define GUDA_FUSED_MAX_OPS 8
#define GUDA_FUSED_MAX_OTHERS 4
typedef struct { int type; float alpha; long long broadcast_size; long long broadcast_dim; } guda_fused_op_t;
int guda_fused_execute(float* x, const float* const* others, int n_others, long long n, const guda_fused_op_t* ops, int n_ops,
                       guda_stream_t stream);

/* ---- Conv2D: conv.go:9-31 ConvParams, :69 Conv2D = im2col -> Sgemm -> NCHW placement + bias.
 * input [B][IC][H][W], kernel [OC][IC][KH][KW], bias [OC] (used when use_bias != 0 and bias != NULL), output [B][OC][OH][OW].
 * The implicit-GEMM route is stream-ordered (one kernel launch once a geometry's offset table exists); the materialised route
 * blocks until done (it owns a pooled workspace, like the reference's `defer Free(workspace)`). */
typedef struct {
  int batch_size, in_channels, in_height, in_width;
  int out_channels, kernel_height, kernel_width;
  int stride_h, stride_w, pad_h, pad_w, dilation_h, dilation_w;
  int use_bias;
} guda_conv_params_t;
int guda_conv2d(const float* input, const float* kernel, const float* bias, float* output, const guda_conv_params_t* params,
                guda_stream_t stream);
/* conv.go:231-300 Conv2DDirect: the reference's direct loops (unfused multiply-add over the in-bounds taps in (ic, kh, kw) order, no
 * zero test, then + bias), bit-identical; stream-ordered.  The parity cross-check API -- Conv2D is the fast route. */
int guda_conv2d_direct(const float* input, const float* kernel, const float* bias, float* output, const guda_conv_params_t* params,
                       guda_stream_t stream);
/* Two routes with the same FMA chain per output element (bit-identical results): 0 = implicit GEMM (default: the column matrix of
 * conv.go:200-222 is gathered inside the GEMM's tile loader, bias + NCHW placement fused in its epilogue), 1 = materialised
 * im2col -> guda_sgemm -> placement (also taken below 4096 output positions, for the 4^3/8^3/16^3 Sgemm fast-path shapes and in
 * the 3xTF32 mode);
 * 2 / 3 / 4 = implicit GEMM with the 64x128 / 64x64 / 128x128 CTA tile forced (the default picks by estimated makespan);
 * -1 = back to the default / GUDA_CONV_ROUTE=im2col environment override.  For A/B measurement and the route-equality tests. */
int guda_conv_set_route(int route);

/* ---- conversions: bfloat16.go:12-29, float16.go:21-87 (device-side bulk convert) ---- */
int guda_f32_to_bf16(const float* src, uint16_t* dst, size_t n, guda_stream_t stream);
int guda_bf16_to_f32(const uint16_t* src, float* dst, size_t n, guda_stream_t stream);
int guda_f32_to_f16(const float* src, uint16_t* dst, size_t n, guda_stream_t stream);
int guda_f16_to_f32(const uint16_t* src, float* dst, size_t n, guda_stream_t stream);

/* ---- synthetic data on the device (bench / large-config parity): counter-based, any sub-block reproducible ---- */
int guda_fill_uniform_f32(float* dst, size_t rows, size_t cols, size_t ld, uint64_t seed, uint64_t row0, uint64_t col0,
                          uint64_t full_cols, float lo, float hi, guda_stream_t stream);
int guda_fill_uniform_f64(double* dst, size_t rows, size_t cols, size_t ld, uint64_t seed, uint64_t row0, uint64_t col0,
                          uint64_t full_cols, double lo, double hi, guda_stream_t stream);

/* ---- measured denominators for rooflines (bench.py): kind 0 = FP32 FFMA TFLOP/s, 1 = FP64 DFMA TFLOP/s,
 * 2 = 128-bit streaming copy GB/s (read+write bytes), 3 = FP64 tensor pipe (DMMA m8n8k4) TFLOP/s,
 * 4 = packed fp32 FMA (fma.rn.f32x2 / FFMA2) TFLOP/s ---- */
/* ---- tolerance framework (tolerance.go:10-191, tolerance_arch.go:8-127, numerical_parity.go:103-141) ------------------------------
 * ToleranceConfig with the additions the "sm_100a" architecture entry needs (zero = not applicable / keep the base value):
 *   k_scale       an element also passes if |delta| <= k_scale * K * u * sum_l |a_il||b_lj| (u = unit roundoff of the accumulation type)
 *   half_ulp_tol  bf16 / fp16 outputs: distance in ULPs of the 16-bit format
 *   abs64, rel64  fp64 results (blas/testblas/common.go:22-38 dTolEqual) */
typedef struct guda_tolerance {
  float abs_tol, rel_tol;
  int ulp_tol;
  int check_nan, check_inf;
  float k_scale;
  int half_ulp_tol;
  double abs64, rel64;
} guda_tolerance_t;
/* tolerance.go:121-128 VerificationResult */
typedef struct guda_verify_result {
  float max_abs_error, max_rel_error;
  int max_ulp_error;
  long long num_errors, total_items, first_error;      /* first_error = -1 if none */
} guda_verify_result_t;
/* 0 DefaultTolerance, 1 StrictTolerance, 2 RelaxedTolerance (tolerance.go:28-54) */
int guda_tolerance_preset(int preset, guda_tolerance_t* out);
/* tolerance_arch.go:41-57 mergeTolerances: non-zero override fields replace the base */
int guda_merge_tolerances(const guda_tolerance_t* base, const guda_tolerance_t* override_, guda_tolerance_t* out);
/* tolerance_arch.go:116-127 GetOperationTolerance.  operation: "gemm", "conv2d", "reduce_sum", "softmax" as in the reference, plus
 * "dot", "dgemm", "gemm_3xtf32", "gemm_bf16", "gemm_f16", "gemm_bf16_fused"; anything else -> DefaultTolerance.
 * arch: NULL or "sm_100a" (the device that produced the results), "amd64", "arm64", "generic" (the reference's GOARCH cases). */
int guda_get_operation_tolerance(const char* operation, const char* arch, guda_tolerance_t* out);
/* numerical_parity.go:132-141 StandardTolerances[operation]; NotImplemented for a name the map does not hold */
int guda_standard_tolerance(const char* operation, float* abs_tol, float* rel_tol, int* ulp_tol);
/* tolerance.go:131-174 VerifyFloat32Array on two DEVICE arrays (one HBM-streaming pass); blocks, result in host memory */
int guda_verify_f32(const float* expected, const float* actual, long long n, const guda_tolerance_t* tol,
                    guda_verify_result_t* host_result, guda_stream_t stream);

int guda_microbench_peak(int kind, int reps, double* value);

/* ---- multi-GPU (new, additive: GetDeviceCount()==1 stays pinned for the single-device API, guda_test.go:332) ----
 * One process per GPU.  The communicator wraps NCCL (loaded at run time); id bytes come from
 * guda_comm_unique_id on rank 0 and travel out of band (the Go shim or torch.distributed). */
typedef struct guda_comm_s* guda_comm_t;
#define GUDA_COMM_ID_BYTES 128
int guda_comm_unique_id(uint8_t host_id[GUDA_COMM_ID_BYTES]);
int guda_comm_init_rank(guda_comm_t* comm, int nranks, int rank, const uint8_t host_id[GUDA_COMM_ID_BYTES]);
int guda_comm_destroy(guda_comm_t comm);
/* 2-D block-partitioned SGEMM: this rank owns A row panel (m_local x k), B column panel (k x n_local), computes
 * its C block with full K locally, then all-gathers the blocks so c_full (m_total x n_total, ld = ldc_full) is
 * complete on every rank.  grid = pr x pc, rank = r*pc + c. */
int guda_sgemm_sharded(guda_comm_t comm, int pr, int pc, int m_total, int n_total, int k, float alpha,
                       const float* a_panel, int lda, const float* b_panel, int ldb, float* c_block, int ldc_block,
                       float* c_full, int ldc_full, guda_stream_t stream);
/* The same step from pinned HOST panels to a HOST C block; a_panel_dev (m/pr x k), b_panel_dev (k x n/pc) and c_block_dev
 * (m/pr x n/pc) are caller-provided device staging.  Uploads are pipelined under the GEMM by 8192-deep K chunks (same bits: each
 * chunk resumes the stored FMA chains with beta = 1); the last chunk is the fused GEMM + all-gather launch.  Blocks. */
int guda_sgemm_sharded_host(guda_comm_t comm, int pr, int pc, int m_total, int n_total, int k, float alpha, const float* host_a_panel, int lda,
                            const float* host_b_panel, int ldb, float* host_c_block, int ldc_host, float* a_panel_dev, float* b_panel_dev,
                            float* c_block_dev, float* c_full, int ldc_full, guda_stream_t stream);
/* per-GPU partial dot products combined by ncclAllReduce(sum) */
/* how guda_sgemm_sharded gathers C: P2P (default) = ONE kernel, the GEMM epilogue stores each finished tile straight into every
 * rank's c_full through NVLink-mapped peer pointers (CUDA IPC), no staging copy, no unpack pass; NCCL = local GEMM, then
 * ncclAllGather (+ unpack).  If peer mapping is unavailable on every rank's agreement the call falls back to NCCL. */
enum { GUDA_GATHER_P2P = 0, GUDA_GATHER_NCCL = 1 };
int guda_comm_set_gather_mode(guda_comm_t comm, int mode);
int guda_comm_get_gather_mode(guda_comm_t comm, int* mode_in_effect);
int guda_sdot_sharded(guda_comm_t comm, int n_local, const float* x, const float* y, float* host_result, guda_stream_t stream);
/* the same, stream-ordered: the sum lands in dev_result (device memory) on every rank and the call does not block */
int guda_sdot_sharded_device(guda_comm_t comm, int n_local, const float* x, const float* y, float* dev_result, guda_stream_t stream);
/* AXPY on this rank's contiguous chunk of a sharded vector (math.go:39): elementwise, no collective. */
int guda_saxpy_sharded(guda_comm_t comm, int n_local, float alpha, const float* x, float* y, guda_stream_t stream);
/* DGEMM twin of guda_sgemm_sharded (compute/dgemm.go:25): 2-D block partition, full K per rank (bit-identical to one GPU),
 * C gathered on every rank -- by the DMMA kernel's peer-store epilogue (default) or ncclAllGather. */
int guda_dgemm_sharded(guda_comm_t comm, int pr, int pc, int m_total, int n_total, int k, double alpha,
                       const double* a_panel, int lda, const double* b_panel, int ldb,
                       double* c_block, int ldc_block, double* c_full, int ldc_full, guda_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif
