"""ctypes loader for the CPU oracle (TEST INFRASTRUCTURE -- never imported by guda_b200).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  It wraps oracle/libguda_oracle.so (built by
oracle/Makefile) with numpy-friendly helpers.  See guda_oracle.h for citations.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libguda_oracle.so")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("level1.c", "gemm.c", "fused.c", "guda_oracle.h", "Makefile")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


class Tol(C.Structure):
    _fields_ = [("abs_tol", C.c_float), ("rel_tol", C.c_float), ("ulp_tol", C.c_int),
                ("check_nan", C.c_int), ("check_inf", C.c_int)]


class Verify(C.Structure):
    _fields_ = [("max_abs", C.c_float), ("max_rel", C.c_float), ("max_ulp", C.c_int),
                ("num_errors", C.c_int64), ("total", C.c_int64), ("first_error", C.c_int64)]


class FusedOp(C.Structure):
    """orc_fused_op_t: type 0 Add, 1 Mul, 2 AddScalar, 3 MulScalar, 4 ReLU, 5 Sigmoid, 6 Custom(geluFloat32)."""
    _fields_ = [("type", C.c_int), ("alpha", C.c_float), ("bsize", C.c_longlong), ("bdim", C.c_longlong)]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        fp, dp, u16p = C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_uint16)
        sz, i, f, d = C.c_size_t, C.c_int, C.c_float, C.c_double
        L.orc_panic_string.restype = C.c_char_p; L.orc_panic_string.argtypes = [i]
        L.orc_saxpy.argtypes = [i, f, fp, sz, i, fp, sz, i]
        L.orc_sdot.argtypes = [i, fp, sz, i, fp, sz, i, fp]
        L.orc_sscal.argtypes = [i, f, fp, sz, i]
        L.orc_dsdot.argtypes = [i, fp, sz, i, fp, sz, i, dp]
        L.orc_sdsdot.argtypes = [i, f, fp, sz, i, fp, sz, i, fp]
        L.orc_daxpy.argtypes = [i, d, dp, sz, i, dp, sz, i]
        L.orc_ddot.argtypes = [i, dp, sz, i, dp, sz, i, dp]
        L.orc_dscal.argtypes = [i, d, dp, sz, i]
        L.orc_sgemm.argtypes = [i, i, i, i, i, f, fp, sz, i, fp, sz, i, f, fp, sz, i, i, i]
        L.orc_dgemm.argtypes = [i, i, i, i, i, d, dp, sz, i, dp, sz, i, d, dp, sz, i]
        L.orc_ref_axpy.argtypes = [f, fp, fp, sz]; L.orc_ref_axpy.restype = None
        L.orc_ref_dot.argtypes = [fp, fp, sz]; L.orc_ref_dot.restype = f
        L.orc_ref_scale.argtypes = [f, fp, sz]; L.orc_ref_scale.restype = None
        L.orc_ref_gemm.argtypes = [i, i, i, i, i, f, fp, i, fp, i, f, fp, i]; L.orc_ref_gemm.restype = None
        L.orc_ref_gelu.argtypes = [fp, sz]; L.orc_ref_gelu.restype = None
        for name in ("orc_exp_f32", "orc_tanh_f32", "orc_gelu_f32", "orc_gelu_pade"):
            getattr(L, name).argtypes = [f]; getattr(L, name).restype = f
        L.orc_fused_gemm_bias_relu.argtypes = [i, i, i, i, i, f, fp, i, fp, i, f, fp, i, fp, i]
        L.orc_fused_gemm_bias_gelu.argtypes = [i, i, i, i, i, f, fp, i, fp, i, f, fp, i, fp, i, i]
        L.orc_add_bias_gelu.argtypes = [fp, fp, i, i]; L.orc_add_bias_gelu.restype = None
        L.orc_sigmoid_f32.argtypes = [f]; L.orc_sigmoid_f32.restype = f
        L.orc_ew_add.argtypes = [fp, fp, fp, sz]; L.orc_ew_add.restype = None
        L.orc_ew_mul.argtypes = [fp, fp, fp, sz]; L.orc_ew_mul.restype = None
        L.orc_ew_relu.argtypes = [fp, sz]; L.orc_ew_relu.restype = None
        L.orc_ew_sigmoid.argtypes = [fp, sz]; L.orc_ew_sigmoid.restype = None
        L.orc_reduce_axis.argtypes = [fp, i, i, i, i, fp]
        L.orc_topk.argtypes = [fp, C.c_longlong, i, fp, C.POINTER(C.c_longlong)]
        L.orc_fused_execute.argtypes = [fp, C.POINTER(fp), i, sz, C.POINTER(FusedOp), i]; L.orc_fused_execute.restype = None
        L.orc_to_bf16.argtypes = [f]; L.orc_to_bf16.restype = C.c_uint16
        L.orc_bf16_to_f32.argtypes = [C.c_uint16]; L.orc_bf16_to_f32.restype = f
        L.orc_f32_to_f16.argtypes = [f]; L.orc_f32_to_f16.restype = C.c_uint16
        L.orc_f16_to_f32.argtypes = [C.c_uint16]; L.orc_f16_to_f32.restype = f
        L.orc_gemm_bf16.argtypes = [i, i, i, f, u16p, i, u16p, i, f, u16p, i]; L.orc_gemm_bf16.restype = None
        L.orc_gemm_f16.argtypes = [i, i, i, i, i, f, u16p, i, u16p, i, f, u16p, i, i]
        L.orc_sgemv.argtypes = [i, i, i, f, fp, sz, i, fp, sz, i, f, fp, sz, i]
        L.orc_f32_gemv.argtypes = [i, i, i, f, fp, i, fp, i, f, fp, sz, i]; L.orc_f32_gemv.restype = None
        L.orc_dgemv.argtypes = [i, i, i, d, dp, sz, i, dp, sz, i, d, dp, sz, i]
        L.orc_dger.argtypes = [i, i, d, dp, sz, i, dp, sz, i, dp, sz, i]
        L.orc_sger.argtypes = [i, i, f, fp, sz, i, fp, sz, i, fp, sz, i]
        for name in ("orc_ssum", "orc_ssumsq", "orc_smax", "orc_smin"):
            getattr(L, name).argtypes = [fp, sz]; getattr(L, name).restype = f
        for name in ("orc_sargmax", "orc_sargmin"):
            getattr(L, name).argtypes = [fp, sz]; getattr(L, name).restype = C.c_longlong
        L.orc_fast_exp.argtypes = [f]; L.orc_fast_exp.restype = f
        L.orc_softmax.argtypes = [fp, sz]; L.orc_softmax.restype = None
        L.orc_layer_norm.argtypes = [fp, fp, fp, sz, f]; L.orc_layer_norm.restype = None
        L.orc_isamax.argtypes = [i, fp, sz, i, C.POINTER(C.c_longlong)]
        L.orc_generate_f32.argtypes = [fp, sz, C.c_uint64]; L.orc_generate_f32.restype = None
        L.orc_generate_f32_masked.argtypes = [fp, sz, C.c_uint64, f, f]; L.orc_generate_f32_masked.restype = None
        L.orc_generate_f64_masked.argtypes = [dp, sz, C.c_uint64, d, d]; L.orc_generate_f64_masked.restype = None
        L.orc_f32_ulp_diff.argtypes = [f, f]; L.orc_f32_ulp_diff.restype = i
        L.orc_f32_near_equal.argtypes = [f, f, C.POINTER(Tol)]; L.orc_f32_near_equal.restype = i
        L.orc_verify_f32.argtypes = [fp, fp, sz, C.POINTER(Tol), C.POINTER(Verify)]; L.orc_verify_f32.restype = None
        _lib = L
    return _lib


class OraclePanic(Exception):
    """Mirrors a Go panic(string) from compute/errors.go."""


def _chk(rc):
    if rc:
        raise OraclePanic(lib().orc_panic_string(rc).decode())


def _p(a, ct):
    return a.ctypes.data_as(C.POINTER(ct))


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a


# ---------------------------------------------------------------- Level 1
def saxpy(n, alpha, x, incx, y, incy):
    """In place on y (float32 array)."""
    assert y.dtype == np.float32 and y.flags.c_contiguous
    x = _f32(x)
    _chk(lib().orc_saxpy(n, alpha, _p(x, C.c_float), x.size, incx, _p(y, C.c_float), y.size, incy))
    return y


def sdot(n, x, incx, y, incy):
    x, y = _f32(x), _f32(y)
    r = C.c_float(0)
    _chk(lib().orc_sdot(n, _p(x, C.c_float), x.size, incx, _p(y, C.c_float), y.size, incy, C.byref(r)))
    return np.float32(r.value)


def sscal(n, alpha, x, incx):
    assert x.dtype == np.float32 and x.flags.c_contiguous
    _chk(lib().orc_sscal(n, alpha, _p(x, C.c_float), x.size, incx))
    return x


def daxpy(n, alpha, x, incx, y, incy):
    assert y.dtype == np.float64 and y.flags.c_contiguous
    x = np.ascontiguousarray(x, dtype=np.float64)
    _chk(lib().orc_daxpy(n, alpha, _p(x, C.c_double), x.size, incx, _p(y, C.c_double), y.size, incy))
    return y


def ddot(n, x, incx, y, incy):
    x = np.ascontiguousarray(x, dtype=np.float64); y = np.ascontiguousarray(y, dtype=np.float64)
    r = C.c_double(0)
    _chk(lib().orc_ddot(n, _p(x, C.c_double), x.size, incx, _p(y, C.c_double), y.size, incy, C.byref(r)))
    return r.value


def dscal(n, alpha, x, incx):
    assert x.dtype == np.float64 and x.flags.c_contiguous
    _chk(lib().orc_dscal(n, alpha, _p(x, C.c_double), x.size, incx))
    return x


# ---------------------------------------------------------------- Level 3
def sgemm(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stock=False, threads=1):
    """In place on c (flat float32 array)."""
    assert c.dtype == np.float32 and c.flags.c_contiguous
    a, b = _f32(a), _f32(b)
    _chk(lib().orc_sgemm(int(ta), int(tb), m, n, k, alpha, _p(a, C.c_float), a.size, lda,
                         _p(b, C.c_float), b.size, ldb, beta, _p(c, C.c_float), c.size, ldc, int(stock), threads))
    return c


def dgemm(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc):
    assert c.dtype == np.float64 and c.flags.c_contiguous
    a = np.ascontiguousarray(a, dtype=np.float64); b = np.ascontiguousarray(b, dtype=np.float64)
    _chk(lib().orc_dgemm(int(ta), int(tb), m, n, k, alpha, _p(a, C.c_double), a.size, lda,
                         _p(b, C.c_double), b.size, ldb, beta, _p(c, C.c_double), c.size, ldc))
    return c


def ref_gemm(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc):
    a, b = _f32(a), _f32(b)
    lib().orc_ref_gemm(int(ta), int(tb), m, n, k, alpha, _p(a, C.c_float), lda, _p(b, C.c_float), ldb,
                       beta, _p(c, C.c_float), ldc)
    return c


def ref_dot(x, y):
    x, y = _f32(x), _f32(y)
    return np.float32(lib().orc_ref_dot(_p(x, C.c_float), _p(y, C.c_float), x.size))


def ref_axpy(alpha, x, y):
    x = _f32(x)
    lib().orc_ref_axpy(alpha, _p(x, C.c_float), _p(y, C.c_float), x.size)
    return y


# ---------------------------------------------------------------- activations / fused
def gelu(x, route="scalar"):
    fn = lib().orc_gelu_f32 if route == "scalar" else lib().orc_gelu_pade
    x = _f32(x)
    return np.array([fn(float(v)) for v in x.ravel()], dtype=np.float32).reshape(x.shape)


def tanh(x):
    x = _f32(x)
    return np.array([lib().orc_tanh_f32(float(v)) for v in x.ravel()], dtype=np.float32).reshape(x.shape)


def exp(x):
    x = _f32(x)
    return np.array([lib().orc_exp_f32(float(v)) for v in x.ravel()], dtype=np.float32).reshape(x.shape)


def ew_add(a, b):                       # math.go:133
    a, b = _f32(a), _f32(b); c = np.empty_like(a)
    lib().orc_ew_add(_p(a, C.c_float), _p(b, C.c_float), _p(c, C.c_float), a.size)
    return c


def ew_mul(a, b):                       # math.go:152
    a, b = _f32(a), _f32(b); c = np.empty_like(a)
    lib().orc_ew_mul(_p(a, C.c_float), _p(b, C.c_float), _p(c, C.c_float), a.size)
    return c


def ew_relu(x):                         # math.go:236 (in place on a copy)
    x = _f32(x).copy(); lib().orc_ew_relu(_p(x, C.c_float), x.size); return x


def ew_sigmoid(x):                      # math.go:255 -> activations.go:9
    x = _f32(x).copy(); lib().orc_ew_sigmoid(_p(x, C.c_float), x.size); return x


def reduce_axis(x, m, n, axis, is_max=False):          # reduce.go:105 ReduceSum / :139 ReduceMax
    x = _f32(x)
    out = np.empty(n if axis == 0 else m, np.float32)
    _chk(lib().orc_reduce_axis(_p(x, C.c_float), m, n, axis, int(is_max), _p(out, C.c_float)))
    return out


def topk(x, k):                         # reduce.go:247 TopK -> (values, indices)
    x = _f32(x)
    kk = min(k, x.size)
    vals, idx = np.empty(max(kk, 1), np.float32), np.empty(max(kk, 1), np.int64)
    got = lib().orc_topk(_p(x, C.c_float), x.size, k, _p(vals, C.c_float), _p(idx, C.c_longlong))
    if got < 0:
        raise ValueError("invalid parameter")
    return vals[:got], idx[:got]


def fused_execute(x, others, ops):
    """fused.go:123 FusedKernel.Execute on a copy of x.  ops: list of (type, alpha, bsize, bdim)."""
    x = _f32(x).copy()
    others = [_f32(o) for o in others]
    fp = C.POINTER(C.c_float)
    arr = (fp * max(len(others), 1))(*[_p(o, C.c_float) for o in others])
    oparr = (FusedOp * max(len(ops), 1))(*[FusedOp(int(t), float(a), int(bs), int(bd)) for t, a, bs, bd in ops])
    lib().orc_fused_execute(_p(x, C.c_float), arr, len(others), x.size, oparr, len(ops))
    return x


def ref_gelu(x):
    x = _f32(x).copy()
    lib().orc_ref_gelu(_p(x, C.c_float), x.size)
    return x


def add_bias_gelu(x, bias):
    bias = _f32(bias)
    lib().orc_add_bias_gelu(_p(x, C.c_float), _p(bias, C.c_float), x.size, bias.size)
    return x


def fused_gemm_bias_relu(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, threads=1):
    a, b, bias = _f32(a), _f32(b), _f32(bias)
    _chk(lib().orc_fused_gemm_bias_relu(int(ta), int(tb), m, n, k, alpha, _p(a, C.c_float), lda, _p(b, C.c_float), ldb,
                                        beta, _p(c, C.c_float), ldc, _p(bias, C.c_float), threads))
    return c


def fused_gemm_bias_gelu(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, route="scalar", threads=1):
    a, b, bias = _f32(a), _f32(b), _f32(bias)
    _chk(lib().orc_fused_gemm_bias_gelu(int(ta), int(tb), m, n, k, alpha, _p(a, C.c_float), lda, _p(b, C.c_float), ldb,
                                        beta, _p(c, C.c_float), ldc, _p(bias, C.c_float),
                                        0 if route == "scalar" else 1, threads))
    return c


def dsdot(n, x, incx, y, incy):
    x, y = _f32(x), _f32(y)
    r = C.c_double(0)
    _chk(lib().orc_dsdot(n, _p(x, C.c_float), x.size, incx, _p(y, C.c_float), y.size, incy, C.byref(r)))
    return r.value


def sdsdot(n, alpha, x, incx, y, incy):
    x, y = _f32(x), _f32(y)
    r = C.c_float(0)
    _chk(lib().orc_sdsdot(n, C.c_float(alpha), _p(x, C.c_float), x.size, incx, _p(y, C.c_float), y.size, incy, C.byref(r)))
    return np.float32(r.value)


# ---------------------------------------------------------------- Level 2
def sgemv(trans, m, n, alpha, a, lda, x, incx, beta, y, incy):
    """In place on y (float32 array)."""
    assert y.dtype == np.float32 and y.flags.c_contiguous
    a, x = _f32(a), _f32(x)
    _chk(lib().orc_sgemv(int(trans), m, n, C.c_float(alpha), _p(a, C.c_float), C.c_size_t(a.size), lda, _p(x, C.c_float), C.c_size_t(x.size), incx,
                         C.c_float(beta), _p(y, C.c_float), C.c_size_t(y.size), incy))
    return y


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def dgemv(trans, m, n, alpha, a, lda, x, incx, beta, y, incy):
    """In place on y (float64 array)."""
    assert y.dtype == np.float64 and y.flags.c_contiguous
    a, x = _f64(a), _f64(x)
    _chk(lib().orc_dgemv(int(trans), m, n, C.c_double(alpha), _p(a, C.c_double), C.c_size_t(a.size), lda, _p(x, C.c_double), C.c_size_t(x.size), incx,
                         C.c_double(beta), _p(y, C.c_double), C.c_size_t(y.size), incy))
    return y


def dger(m, n, alpha, x, incx, y, incy, a, lda):
    """In place on a (float64 array)."""
    assert a.dtype == np.float64 and a.flags.c_contiguous
    x, y = _f64(x), _f64(y)
    _chk(lib().orc_dger(m, n, C.c_double(alpha), _p(x, C.c_double), C.c_size_t(x.size), incx, _p(y, C.c_double), C.c_size_t(y.size), incy,
                        _p(a, C.c_double), C.c_size_t(a.size), lda))
    return a


def f32_gemv(trans, m, n, alpha, a, lda, x, incx, beta, y, incy):
    """package f32's GemvN / GemvT (what gemv_test.go calls): unchecked, no alpha == 0 shortcut.  In place on y."""
    assert y.dtype == np.float32 and y.flags.c_contiguous
    a, x = _f32(a), _f32(x)
    lib().orc_f32_gemv(int(trans), m, n, C.c_float(alpha), _p(a, C.c_float), lda, _p(x, C.c_float), incx, C.c_float(beta),
                       _p(y, C.c_float), C.c_size_t(y.size), incy)
    return y


def sger(m, n, alpha, x, incx, y, incy, a, lda):
    """In place on a (float32 array)."""
    assert a.dtype == np.float32 and a.flags.c_contiguous
    x, y = _f32(x), _f32(y)
    _chk(lib().orc_sger(m, n, C.c_float(alpha), _p(x, C.c_float), C.c_size_t(x.size), incx, _p(y, C.c_float), C.c_size_t(y.size), incy,
                        _p(a, C.c_float), C.c_size_t(a.size), lda))
    return a


# ---------------------------------------------------------------- Reductions (reduce.go, compute/asm/f32/reduce*.s, sum_amd64.s)
def _red(name, x, n=None):
    x = _f32(x)
    n = x.size if n is None else n
    return getattr(lib(), name)(_p(x, C.c_float), C.c_size_t(n))


def ssum(x, n=None): return np.float32(_red("orc_ssum", x, n))
def ssumsq(x, n=None): return np.float32(_red("orc_ssumsq", x, n))
def smax(x, n=None): return np.float32(_red("orc_smax", x, n))
def smin(x, n=None): return np.float32(_red("orc_smin", x, n))
def sargmax(x, n=None): return int(_red("orc_sargmax", x, n))
def sargmin(x, n=None): return int(_red("orc_sargmin", x, n))


def fast_exp(x): return np.float32(lib().orc_fast_exp(C.c_float(x)))


def softmax(x):
    """In place on x (float32 array)."""
    assert x.dtype == np.float32 and x.flags.c_contiguous
    lib().orc_softmax(_p(x, C.c_float), C.c_size_t(x.size))
    return x


def layer_norm(x, gamma, beta, eps):
    assert x.dtype == np.float32 and x.flags.c_contiguous
    gp = _p(_f32(gamma), C.c_float) if gamma is not None else None
    bp = _p(_f32(beta), C.c_float) if beta is not None else None
    lib().orc_layer_norm(_p(x, C.c_float), gp, bp, C.c_size_t(x.size), C.c_float(eps))
    return x


def isamax(n, x, incx):
    x = _f32(x)
    r = C.c_longlong(0)
    _chk(lib().orc_isamax(n, _p(x, C.c_float), C.c_size_t(x.size), incx, C.byref(r)))
    return r.value


# ---------------------------------------------------------------- Conv2D (conv.go)
CONV_KEYS = ("BatchSize", "InChannels", "InHeight", "InWidth", "OutChannels", "KernelHeight", "KernelWidth",
             "StrideH", "StrideW", "PadH", "PadW", "DilationH", "DilationW")


def conv_out_hw(p):
    """ConvParams.OutputHeight/OutputWidth (conv.go:54-63)."""
    if p["StrideH"] <= 0 or p["StrideW"] <= 0:
        return 0, 0
    oh = (p["InHeight"] + 2 * p["PadH"] - ((p["KernelHeight"] - 1) * p["DilationH"] + 1)) // p["StrideH"] + 1
    ow = (p["InWidth"] + 2 * p["PadW"] - ((p["KernelWidth"] - 1) * p["DilationW"] + 1)) // p["StrideW"] + 1
    return oh, ow


def conv2d(inp, kernel, bias, p, direct=False, threads=1):
    """p: dict with the ConvParams field names.  Returns the NCHW output as a flat float32 array."""
    inp, kernel = _f32(inp), _f32(kernel)
    oh, ow = conv_out_hw(p)
    out = np.zeros(max(p["BatchSize"] * p["OutChannels"] * oh * ow, 1), dtype=np.float32)
    use_bias = int(bool(p.get("UseBias")) and bias is not None)
    b = _f32(bias) if bias is not None else np.zeros(1, dtype=np.float32)
    args = [_p(inp, C.c_float), _p(kernel, C.c_float), _p(b, C.c_float), _p(out, C.c_float)] + [int(p[k]) for k in CONV_KEYS] + [use_bias]
    rc = lib().orc_conv2d_direct(*args) if direct else lib().orc_conv2d(*(args + [threads]))
    if rc:
        raise OraclePanic("conv2d: invalid parameters (code %d)" % rc)
    return out[:p["BatchSize"] * p["OutChannels"] * oh * ow]


# ---------------------------------------------------------------- bf16 / fp16
def to_bf16(x):
    x = _f32(x)
    return np.array([lib().orc_to_bf16(float(v)) for v in x.ravel()], dtype=np.uint16).reshape(x.shape)


def to_bf16_fast(x):
    """Vectorised numpy twin of orc_to_bf16 (checked against it in the tests)."""
    bits = _f32(x).view(np.uint32).astype(np.uint64)
    bump = ((bits & 0x7FFF) == 0x7FFF) & (((bits >> 16) & 1) == 1)
    bits = np.where(bump, bits + 0x10000, bits)
    return ((bits >> 16) & 0xFFFF).astype(np.uint16)


def bf16_to_f32(b):
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)


def f32_to_f16(x):
    x = _f32(x)
    return np.array([lib().orc_f32_to_f16(float(v)) for v in x.ravel()], dtype=np.uint16).reshape(x.shape)


def f16_to_f32(h):
    h = np.ascontiguousarray(h, dtype=np.uint16)
    return np.array([lib().orc_f16_to_f32(int(v)) for v in h.ravel()], dtype=np.float32).reshape(h.shape)


def gemm_bf16(m, n, k, alpha, a, lda, b, ldb, beta, c, ldc):
    a = np.ascontiguousarray(a, dtype=np.uint16); b = np.ascontiguousarray(b, dtype=np.uint16)
    assert c.dtype == np.uint16
    lib().orc_gemm_bf16(m, n, k, alpha, _p(a, C.c_uint16), lda, _p(b, C.c_uint16), ldb, beta, _p(c, C.c_uint16), ldc)
    return c


def gemm_f16(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, threads=1):
    a = np.ascontiguousarray(a, dtype=np.uint16); b = np.ascontiguousarray(b, dtype=np.uint16)
    assert c.dtype == np.uint16
    _chk(lib().orc_gemm_f16(int(ta), int(tb), m, n, k, alpha, _p(a, C.c_uint16), lda, _p(b, C.c_uint16), ldb,
                            beta, _p(c, C.c_uint16), ldc, threads))
    return c


# ---------------------------------------------------------------- generators / tolerance
def generate_f32(n, seed):
    out = np.empty(n, dtype=np.float32)
    lib().orc_generate_f32(_p(out, C.c_float), n, seed)
    return out


def uniform_f32(n, seed, lo=-1.0, hi=1.0):
    out = np.empty(n, dtype=np.float32)
    lib().orc_generate_f32_masked(_p(out, C.c_float), n, seed, lo, hi)
    return out


def uniform_f64(n, seed, lo=-1.0, hi=1.0):
    out = np.empty(n, dtype=np.float64)
    lib().orc_generate_f64_masked(_p(out, C.c_double), n, seed, lo, hi)
    return out


def verify_f32(expected, actual, tol):
    e, a = _f32(expected).ravel(), _f32(actual).ravel()
    assert e.size == a.size
    t = Tol(tol["abs"], tol["rel"], tol["ulp"], 1, 1)
    v = Verify()
    lib().orc_verify_f32(_p(e, C.c_float), _p(a, C.c_float), e.size, C.byref(t), C.byref(v))
    return {"max_abs": v.max_abs, "max_rel": v.max_rel, "max_ulp": v.max_ulp, "num_errors": v.num_errors,
            "total": v.total, "first_error": v.first_error}


def near_equal(a, b, tol):
    t = Tol(tol["abs"], tol["rel"], tol["ulp"], 1, 1)
    return bool(lib().orc_f32_near_equal(a, b, C.byref(t)))
