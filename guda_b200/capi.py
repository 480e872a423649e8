"""ctypes binding of include/guda_b200.h -- the same C ABI the Go cgo shim binds.

No fallback of any kind: if the shared library is missing, or a compute entry point is
called without a CUDA device, an exception is raised.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GUDA_B200_LIB: load another build of the same library (A/B timing of kernel variants); the default is the in-tree build
LIB_PATH = os.environ.get("GUDA_B200_LIB") or os.path.join(_HERE, "lib", "libguda_b200.so")

OK, ERR_MEMORY, ERR_INVALID_ARG, ERR_EXECUTION, ERR_NUMERICAL, ERR_DEVICE, ERR_NOT_IMPLEMENTED = range(7)
ERROR_TYPE_NAMES = {1: "Memory", 2: "InvalidArgument", 3: "Execution", 4: "Numerical", 5: "Device", 6: "NotImplemented"}
MEMCPY_H2H, MEMCPY_H2D, MEMCPY_D2H, MEMCPY_D2D, MEMCPY_DEFAULT = range(5)
COMM_ID_BYTES = 128


class ConvParams(C.Structure):
    """conv.go:9-31 ConvParams, field for field (guda_conv_params_t)."""
    _fields_ = [(n, C.c_int) for n in ("batch_size", "in_channels", "in_height", "in_width", "out_channels", "kernel_height",
                                       "kernel_width", "stride_h", "stride_w", "pad_h", "pad_w", "dilation_h", "dilation_w", "use_bias")]


class FusedOp(C.Structure):
    """guda_fused_op_t (fused.go:25-32 FusedOp flattened: the broadcast shape becomes (size, dim length))."""
    _fields_ = [("type", C.c_int), ("alpha", C.c_float), ("broadcast_size", C.c_longlong), ("broadcast_dim", C.c_longlong)]


FUSED_ADD, FUSED_MUL, FUSED_ADD_SCALAR, FUSED_MUL_SCALAR, FUSED_RELU, FUSED_SIGMOID, FUSED_GELU = range(7)


class Tolerance(C.Structure):
    """guda_tolerance_t: tolerance.go:10-26 ToleranceConfig + the fields of the sm_100a architecture entry."""
    _fields_ = [("abs_tol", C.c_float), ("rel_tol", C.c_float), ("ulp_tol", C.c_int), ("check_nan", C.c_int), ("check_inf", C.c_int),
                ("k_scale", C.c_float), ("half_ulp_tol", C.c_int), ("abs64", C.c_double), ("rel64", C.c_double)]


class VerifyResult(C.Structure):
    """guda_verify_result_t: tolerance.go:121-128 VerificationResult."""
    _fields_ = [("max_abs_error", C.c_float), ("max_rel_error", C.c_float), ("max_ulp_error", C.c_int), ("num_errors", C.c_longlong),
                ("total_items", C.c_longlong), ("first_error", C.c_longlong)]


class DeviceProps(C.Structure):
    _fields_ = [("id", C.c_int), ("name", C.c_char * 128), ("total_mem", C.c_size_t), ("sm_count", C.c_int),
                ("max_threads", C.c_int), ("cc_major", C.c_int), ("cc_minor", C.c_int), ("l2_bytes", C.c_int)]


# name -> (restype, argtypes); every symbol include/guda_b200.h declares
vp, i, f, d, sz, u64 = C.c_void_p, C.c_int, C.c_float, C.c_double, C.c_size_t, C.c_uint64
ip, fp, dp, szp, i64p = C.POINTER(C.c_int), C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_size_t), C.POINTER(C.c_int64)
SIGNATURES = {
    "guda_last_error": (C.c_char_p, []),
    "guda_version": (C.c_char_p, []),
    "guda_init": (i, [i]),
    "guda_device_count": (i, [ip]),
    "guda_set_device": (i, [i]),
    "guda_get_device": (i, [ip]),
    "guda_device_properties": (i, [i, C.POINTER(DeviceProps)]),
    "guda_device_synchronize": (i, []),
    "guda_mem_get_info": (i, [szp, szp]),
    "guda_malloc": (i, [C.POINTER(vp), sz]),
    "guda_free": (i, [vp]),
    "guda_pool_stats": (i, [i64p, i64p]),
    "guda_pool_trim": (i, []),
    "guda_malloc_host": (i, [C.POINTER(vp), sz]),
    "guda_free_host": (i, [vp]),
    "guda_memcpy": (i, [vp, vp, sz, i, vp]),
    "guda_memcpy_async": (i, [vp, vp, sz, i, vp]),
    "guda_memset": (i, [vp, i, sz, vp]),
    "guda_stream_create": (i, [C.POINTER(vp)]),
    "guda_stream_destroy": (i, [vp]),
    "guda_stream_synchronize": (i, [vp]),
    "guda_stream_query": (i, [vp, ip]),
    "guda_stream_wait_event": (i, [vp, vp]),
    "guda_stream_native": (i, [vp, C.POINTER(vp)]),
    "guda_event_create": (i, [C.POINTER(vp)]),
    "guda_event_destroy": (i, [vp]),
    "guda_event_record": (i, [vp, vp]),
    "guda_event_synchronize": (i, [vp]),
    "guda_event_query": (i, [vp, ip]),
    "guda_event_elapsed_ms": (i, [vp, vp, fp]),
    "guda_saxpy": (i, [i, f, vp, i, vp, i, vp]),
    "guda_sscal": (i, [i, f, vp, i, vp]),
    "guda_sdot": (i, [i, vp, i, vp, i, fp, vp]),
    "guda_sdot_device": (i, [i, vp, i, vp, i, vp, vp]),
    "guda_sdot_device_static": (i, [i, vp, i, vp, i, vp, vp]),     # no allocation on the call path (no reference counterpart)
    "guda_daxpy": (i, [i, d, vp, i, vp, i, vp]),
    "guda_dscal": (i, [i, d, vp, i, vp]),
    "guda_ddot": (i, [i, vp, i, vp, i, dp, vp]),
    "guda_ddot_device": (i, [i, vp, i, vp, i, vp, vp]),
    "guda_set_sgemm_mode": (i, [i]),
    "guda_set_sgemm_pipe": (i, [i, i]),
    "guda_get_sgemm_mode": (i, [ip]),
    "guda_sgemm": (i, [i, i, i, i, i, f, vp, i, vp, i, f, vp, i, vp]),
    "guda_sgemm_host": (i, [i, i, i, i, i, f, vp, i, vp, i, f, vp, i, vp]),
    "guda_host_gemm_schedule": (i, [i, i, vp, i, vp, vp]),
    "guda_dgemm": (i, [i, i, i, i, i, d, vp, i, vp, i, d, vp, i, vp]),
    "guda_sgemm_scatter": (i, [i, i, i, i, i, f, vp, i, vp, i, f, vp, i, C.POINTER(vp), i, i, i, i, vp]),
    "guda_dgemm_scatter": (i, [i, i, i, i, i, d, vp, i, vp, i, d, vp, i, C.POINTER(vp), i, i, i, i, vp]),
    "guda_set_dgemm_mode": (i, [i]),
    "guda_get_dgemm_mode": (i, [ip]),
    "guda_gemm_bf16": (i, [i, i, i, i, i, f, vp, i, vp, i, f, vp, i, vp]),
    "guda_gemm_f16": (i, [i, i, i, i, i, f, vp, i, vp, i, f, vp, i, vp]),
    "guda_set_gemm16_mode": (i, [i]),
    "guda_get_gemm16_mode": (i, [ip]),
    "guda_sgemm_bias_relu": (i, [i, i, i, i, i, f, vp, i, vp, i, f, vp, i, vp, vp]),
    "guda_sgemm_bias_gelu": (i, [i, i, i, i, i, f, vp, i, vp, i, f, vp, i, vp, vp]),
    "guda_gemm_bf16_bias_relu": (i, [i, i, i, f, vp, i, vp, i, vp, i, vp, vp]),
    "guda_gemm_bf16_bias_gelu": (i, [i, i, i, f, vp, i, vp, i, vp, i, vp, vp]),
    "guda_add_bias_gelu": (i, [vp, vp, i, i, vp]),
    "guda_add_bias_relu": (i, [vp, vp, i, i, vp]),
    "guda_sgemv": (i, [i, i, i, f, vp, i, vp, i, f, vp, i, vp]),
    "guda_dgemv": (i, [i, i, i, d, vp, i, vp, i, d, vp, i, vp]),
    "guda_dger": (i, [i, i, d, vp, i, vp, i, vp, i, vp]),
    "guda_sger": (i, [i, i, f, vp, i, vp, i, vp, i, vp]),
    "guda_dsdot": (i, [i, vp, i, vp, i, dp, vp]),
    "guda_sdsdot": (i, [i, f, vp, i, vp, i, fp, vp]),
    "guda_ssum": (i, [C.c_longlong, vp, fp, vp]),
    "guda_ssumsq": (i, [C.c_longlong, vp, fp, vp]),
    "guda_smax": (i, [C.c_longlong, vp, fp, vp]),
    "guda_smin": (i, [C.c_longlong, vp, fp, vp]),
    "guda_sargmax": (i, [C.c_longlong, vp, C.POINTER(C.c_longlong), vp]),
    "guda_sargmin": (i, [C.c_longlong, vp, C.POINTER(C.c_longlong), vp]),
    "guda_isamax": (i, [i, vp, i, C.POINTER(C.c_longlong), vp]),
    "guda_softmax": (i, [vp, C.c_longlong, vp]),
    "guda_layer_norm": (i, [vp, vp, vp, C.c_longlong, f, vp]),
    "guda_conv2d": (i, [vp, vp, vp, vp, C.POINTER(ConvParams), vp]),
    "guda_topk": (i, [vp, C.c_longlong, i, fp, i64p, vp]),
    "guda_reduce_sum_axis": (i, [vp, i, i, i, vp, vp]),
    "guda_reduce_max_axis": (i, [vp, i, i, i, vp, vp]),
    "guda_add": (i, [vp, vp, vp, C.c_longlong, vp]),
    "guda_multiply": (i, [vp, vp, vp, C.c_longlong, vp]),
    "guda_relu": (i, [vp, C.c_longlong, vp]),
    "guda_sigmoid": (i, [vp, C.c_longlong, vp]),
    "guda_selftest_activations": (i, [vp, vp]),
    "guda_fused_execute": (i, [vp, C.POINTER(vp), i, C.c_longlong, C.POINTER(FusedOp), i, vp]),
    "guda_conv_set_route": (i, [i]),
    "guda_conv2d_direct": (i, [vp, vp, vp, vp, C.POINTER(ConvParams), vp]),
    "guda_f32_to_bf16": (i, [vp, vp, sz, vp]),
    "guda_bf16_to_f32": (i, [vp, vp, sz, vp]),
    "guda_f32_to_f16": (i, [vp, vp, sz, vp]),
    "guda_f16_to_f32": (i, [vp, vp, sz, vp]),
    "guda_fill_uniform_f32": (i, [vp, sz, sz, sz, u64, u64, u64, u64, f, f, vp]),
    "guda_fill_uniform_f64": (i, [vp, sz, sz, sz, u64, u64, u64, u64, d, d, vp]),
    "guda_comm_unique_id": (i, [C.POINTER(C.c_uint8)]),
    "guda_comm_init_rank": (i, [C.POINTER(vp), i, i, C.POINTER(C.c_uint8)]),
    "guda_comm_destroy": (i, [vp]),
    "guda_sgemm_sharded": (i, [vp, i, i, i, i, i, f, vp, i, vp, i, vp, i, vp, i, vp]),
    "guda_sgemm_sharded_host": (i, [vp, i, i, i, i, i, f, vp, i, vp, i, vp, i, vp, vp, vp, vp, i, vp]),
    "guda_comm_set_gather_mode": (i, [vp, i]),
    "guda_comm_get_gather_mode": (i, [vp, ip]),
    "guda_sdot_sharded": (i, [vp, i, vp, vp, fp, vp]),
    "guda_sdot_sharded_device": (i, [vp, i, vp, vp, vp, vp]),
    "guda_saxpy_sharded": (i, [vp, i, f, vp, vp, vp]),
    "guda_dgemm_sharded": (i, [vp, i, i, i, i, i, d, vp, i, vp, i, vp, i, vp, i, vp]),
    "guda_tolerance_preset": (i, [i, C.POINTER(Tolerance)]),
    "guda_merge_tolerances": (i, [C.POINTER(Tolerance), C.POINTER(Tolerance), C.POINTER(Tolerance)]),
    "guda_get_operation_tolerance": (i, [C.c_char_p, C.c_char_p, C.POINTER(Tolerance)]),
    "guda_standard_tolerance": (i, [C.c_char_p, fp, fp, ip]),
    "guda_verify_f32": (i, [vp, vp, C.c_longlong, C.POINTER(Tolerance), C.POINTER(VerifyResult), vp]),
    "guda_microbench_peak": (i, [i, i, dp]),
}

_lib = None


def load():
    """Load libguda_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(nvcc, sm_100a). There is no CPU fallback." % LIB_PATH)
        lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in SIGNATURES.items():
            if not hasattr(lib, name) and os.environ.get("GUDA_B200_LIB_PARTIAL") == "1":
                continue                     # A/B timing against an older build of the library (scratch/build_variant.sh)
            fn = getattr(lib, name)          # AttributeError if the header and the library disagree
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


class GUDAError(Exception):
    """errors.go:27-44 GUDAError{Type, Op, Message}."""

    def __init__(self, status, op, message):
        self.type, self.op, self.message = status, op, message
        super().__init__("GUDA %s error in %s: %s" % (ERROR_TYPE_NAMES.get(status, "Unknown"), op, message))


class BlasPanic(Exception):
    """A Go panic(string) of compute/errors.go:9-35, e.g. 'blas: m < 0'."""


def check(status, op):
    if status == OK:
        return
    msg = load().guda_last_error().decode()
    if status == ERR_INVALID_ARG and msg.startswith("blas: "):
        raise BlasPanic(msg)
    raise GUDAError(status, op, msg)
