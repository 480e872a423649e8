"""Host-side mirror of GUDA's root-package API (package guda) above the C ABI.

Same names, argument meaning and error behaviour as the reference's Go functions, so the
parity tests read like the reference's own tests:
  Malloc/Free/Memcpy/DevicePtr   memory.go:57-66,81-145,246-317
  Synchronize/CreateStream       guda.go:242-305
  AXPY/DOT/Scale/GEMM            math.go:39,64,170,98
  FusedGEMMBiasReLU/GELU         fused.go:313, fused_gelu.go:65
  GEMMBFloat16/GEMMFloat16       bfloat16.go:104, float16.go:201
plus the documented aliases (docs/manual/08-blas-api.md): Saxpy/Sdot/Sscal/Sgemm/Dgemm...
The Go shim in go/guda/ is the real drop-in; this module is the same layer for the Python
test/bench harness (there is no Go toolchain in the build image).  Host data are numpy arrays.
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import (BlasPanic, GUDAError, MEMCPY_D2D, MEMCPY_D2H, MEMCPY_DEFAULT, MEMCPY_H2D, MEMCPY_H2H)

MemcpyHostToHost, MemcpyHostToDevice, MemcpyDeviceToHost, MemcpyDeviceToDevice, MemcpyDefault = range(5)


class ErrNotSupported(GUDAError):
    """fused.go:11-17 ErrNotSupported."""


def _L():
    return capi.load()


class DevicePtr:
    """memory.go / guda.go:95-99 DevicePtr{ptr,size,offset}.  Float32()/Float64()/Int32()/Byte() are
    device->host SNAPSHOTS (the reference's were live host views): writes must go through Memcpy."""
    __slots__ = ("ptr", "size", "offset")

    def __init__(self, ptr=0, size=0, offset=0):
        self.ptr, self.size, self.offset = ptr, size, offset

    def Offset(self, nbytes):                      # memory.go:303-309
        return DevicePtr(self.ptr + nbytes, self.size - nbytes, self.offset + nbytes)

    def Size(self):                                # memory.go:312-314
        return self.size

    def _snapshot(self, dtype):
        if not self.ptr:
            return None
        n = self.size // np.dtype(dtype).itemsize
        out = np.empty(n, dtype=dtype)
        if n:
            capi.check(_L().guda_memcpy(out.ctypes.data, self.ptr, n * out.itemsize, MEMCPY_D2H, None), "Memcpy")
        return out

    def Float32(self): return self._snapshot(np.float32)
    def Float64(self): return self._snapshot(np.float64)
    def Int32(self): return self._snapshot(np.int32)
    def Byte(self): return self._snapshot(np.uint8)
    def UInt16(self): return self._snapshot(np.uint16)

    # ---- reductions: reduce.go:13-56 (blocking, scalar to the host like the reference's return values)
    def _slice_check(self, n):
        if n < 0 or n > self.size // 4:
            raise IndexError("slice bounds out of range")          # d.Float32()[:n] in the reference

    def _red_f(self, fn, n):
        self._slice_check(n)
        r = C.c_float(0)
        capi.check(getattr(_L(), fn)(n, self.ptr, C.byref(r), None), fn)
        return np.float32(r.value)

    def _red_i(self, fn, n):
        self._slice_check(n)
        r = C.c_longlong(0)
        capi.check(getattr(_L(), fn)(n, self.ptr, C.byref(r), None), fn)
        return int(r.value)

    def Sum(self, n): return self._red_f("guda_ssum", n)                    # reduce.go:13
    def Max(self, n): return self._red_f("guda_smax", n)                    # reduce.go:18
    def Min(self, n): return self._red_f("guda_smin", n)                    # reduce.go:26
    def ArgMax(self, n): return self._red_i("guda_sargmax", n)              # reduce.go:34
    def ArgMin(self, n): return self._red_i("guda_sargmin", n)              # reduce.go:42
    def SumSquares(self, n): return self._red_f("guda_ssumsq", n)           # reduce.go:51

    def Mean(self, n):                                                       # reduce.go:59-64
        return np.float32(0) if n == 0 else np.float32(self.Sum(n) / np.float32(n))

    def __repr__(self):
        return "DevicePtr(0x%x, size=%d, offset=%d)" % (self.ptr, self.size, self.offset)


class Stream:
    """guda.go:57-62 Stream -> a CUDA stream.  Submit(fn) of the reference (a host closure queue) has no GPU meaning;
    kernels are enqueued by passing stream=... to the BLAS calls."""

    def __init__(self, handle):
        self.handle = handle

    def Synchronize(self):                         # guda.go:297-299
        capi.check(_L().guda_stream_synchronize(self.handle), "StreamSynchronize")

    def Query(self):
        done = C.c_int(0)
        capi.check(_L().guda_stream_query(self.handle, C.byref(done)), "StreamQuery")
        return bool(done.value)

    def WaitEvent(self, event):
        capi.check(_L().guda_stream_wait_event(self.handle, event.handle), "StreamWaitEvent")

    def Destroy(self):
        capi.check(_L().guda_stream_destroy(self.handle), "StreamDestroy")
        self.handle = None

    def native(self):
        out = C.c_void_p()
        capi.check(_L().guda_stream_native(self.handle, C.byref(out)), "StreamNative")
        return out.value or 0


class Event:
    def __init__(self):
        h = C.c_void_p()
        capi.check(_L().guda_event_create(C.byref(h)), "EventCreate")
        self.handle = h.value

    def Record(self, stream=None):
        capi.check(_L().guda_event_record(self.handle, _sh(stream)), "EventRecord")

    def Synchronize(self):
        capi.check(_L().guda_event_synchronize(self.handle), "EventSynchronize")

    def Query(self):
        done = C.c_int(0)
        capi.check(_L().guda_event_query(self.handle, C.byref(done)), "EventQuery")
        return bool(done.value)

    def Destroy(self):
        capi.check(_L().guda_event_destroy(self.handle), "EventDestroy")
        self.handle = None


def ElapsedTime(start, stop):
    ms = C.c_float(0)
    capi.check(_L().guda_event_elapsed_ms(start.handle, stop.handle, C.byref(ms)), "EventElapsedTime")
    return ms.value


def _sh(stream):
    return None if stream is None else stream.handle


# ---------------------------------------------------------------- device / runtime
def Init(device=0):
    capi.check(_L().guda_init(device), "Init")


def GetDeviceCount():
    """guda.go:227-229: the single-device API always reports 1 (pinned by guda_test.go:332-335)."""
    return 1


def PhysicalDeviceCount():
    n = C.c_int(0)
    capi.check(_L().guda_device_count(C.byref(n)), "DeviceCount")
    return n.value


def SetDevice(dev):                                # guda.go:216-221
    if dev != 0:
        raise GUDAError(capi.ERR_INVALID_ARG, "SetDevice", "invalid device ID")


def GetDeviceProperties(dev=0):                    # guda.go:232-237
    if dev != 0:
        raise GUDAError(capi.ERR_INVALID_ARG, "GetDeviceProperties", "invalid device ID: %d" % dev)
    cur = C.c_int(0)
    capi.check(_L().guda_get_device(C.byref(cur)), "GetDevice")
    props = capi.DeviceProps()
    capi.check(_L().guda_device_properties(cur.value, C.byref(props)), "GetDeviceProperties")
    return props


def Synchronize():                                 # guda.go:278-283
    capi.check(_L().guda_device_synchronize(), "Synchronize")


def CreateStream():                                # guda.go:242-256
    h = C.c_void_p()
    capi.check(_L().guda_stream_create(C.byref(h)), "StreamCreate")
    return Stream(h.value)


StreamCreate = CreateStream


def MemGetInfo():
    fr, tot = C.c_size_t(0), C.c_size_t(0)
    capi.check(_L().guda_mem_get_info(C.byref(fr), C.byref(tot)), "MemGetInfo")
    return fr.value, tot.value


# ---------------------------------------------------------------- memory
def Malloc(size):                                  # memory.go:57-59
    p = C.c_void_p()
    capi.check(_L().guda_malloc(C.byref(p), size), "Malloc")
    return DevicePtr(p.value or 0, size, 0)


def Free(ptr):                                     # memory.go:62-66
    capi.check(_L().guda_free(ptr.ptr), "Free")


def GetMemoryStats():                              # memory.go:230-234
    a, p = C.c_int64(0), C.c_int64(0)
    capi.check(_L().guda_pool_stats(C.byref(a), C.byref(p)), "GetStats")
    return a.value, p.value


def PoolTrim():
    """Return cached (freed) blocks to the driver; new, the reference never gives memory back (memory.go:222)."""
    capi.check(_L().guda_pool_trim(), "PoolTrim")


def _addr(obj, writable):
    if isinstance(obj, DevicePtr):
        return obj.ptr
    if isinstance(obj, np.ndarray):
        if not obj.flags.c_contiguous:
            raise GUDAError(capi.ERR_INVALID_ARG, "Memcpy", "host array must be contiguous")
        if writable and not obj.flags.writeable:
            raise GUDAError(capi.ERR_INVALID_ARG, "Memcpy", "destination host array is read-only")
        return obj.ctypes.data if obj.size else 0
    if isinstance(obj, int):
        return obj
    raise GUDAError(capi.ERR_INVALID_ARG, "Memcpy", "unsupported %s type: %s" % ("dst" if writable else "src", type(obj).__name__))


def Memcpy(dst, src, size, kind=MemcpyDefault, stream=None):       # memory.go:81-145
    capi.check(_L().guda_memcpy(_addr(dst, True), _addr(src, False), size, kind, _sh(stream)), "Memcpy")


def MemcpyAsync(dst, src, size, kind=MemcpyDefault, stream=None):
    capi.check(_L().guda_memcpy_async(_addr(dst, True), _addr(src, False), size, kind, _sh(stream)), "MemcpyAsync")


def Memset(ptr, value, size, stream=None):
    capi.check(_L().guda_memset(ptr.ptr, value, size, _sh(stream)), "Memset")


def to_device(arr):
    """Convenience: Malloc + H2D of a numpy array."""
    arr = np.ascontiguousarray(arr)
    d = Malloc(max(arr.nbytes, 1))
    d.size = arr.nbytes
    if arr.nbytes:
        Memcpy(d, arr, arr.nbytes, MemcpyHostToDevice)
    return d


# ---------------------------------------------------------------- slice-length checks done above cgo (DevicePtr knows its size)
def _len(p, itemsize):
    return p.size // itemsize


def _check_level1(n, x, incx, y, incy, item):
    # compute/level1float32.go:226-242 (order and strings)
    if incx == 0: raise BlasPanic("blas: zero x index increment")
    if y is not None and incy == 0: raise BlasPanic("blas: zero y index increment")
    if n < 1:
        if n == 0: return False
        raise BlasPanic("blas: n < 0")
    lx = _len(x, item)
    if (incx > 0 and lx <= (n - 1) * incx) or (incx < 0 and lx <= (1 - n) * incx): raise BlasPanic("blas: insufficient length of x")
    if y is not None:
        ly = _len(y, item)
        if (incy > 0 and ly <= (n - 1) * incy) or (incy < 0 and ly <= (1 - n) * incy): raise BlasPanic("blas: insufficient length of y")
    return True


def _check_gemm_lengths(ta, tb, m, n, k, a, lda, b, ldb, c, ldc, item):
    # compute/sgemm.go:116-136; the dimension/ld checks (:30-71) happen in the C ABI with the same strings
    if m <= 0 or n <= 0 or k < 0: return
    la, lb, lc = _len(a, item), _len(b, item), _len(c, item)
    if min(lda, ldb, ldc) < 1: return
    if (la < (k - 1) * lda + m) if ta else (la < (m - 1) * lda + k): raise BlasPanic("blas: insufficient length of a")
    if (lb < (n - 1) * ldb + k) if tb else (lb < (k - 1) * ldb + n): raise BlasPanic("blas: insufficient length of b")
    if lc < (m - 1) * ldc + n: raise BlasPanic("blas: insufficient length of c")


# ---------------------------------------------------------------- Level 1
def Saxpy(n, alpha, x, incx, y, incy, stream=None):
    if _check_level1(n, x, incx, y, incy, 4):
        capi.check(_L().guda_saxpy(n, alpha, x.ptr, incx, y.ptr, incy, _sh(stream)), "Saxpy")


def Daxpy(n, alpha, x, incx, y, incy, stream=None):
    if _check_level1(n, x, incx, y, incy, 8):
        capi.check(_L().guda_daxpy(n, alpha, x.ptr, incx, y.ptr, incy, _sh(stream)), "Daxpy")


def Sdot(n, x, incx, y, incy, stream=None):
    if incx == 0: raise BlasPanic("blas: zero x index increment")
    if incy == 0: raise BlasPanic("blas: zero y index increment")
    if n <= 0:
        if n == 0: return np.float32(0)
        raise BlasPanic("blas: n < 0")
    _check_level1(n, x, incx, y, incy, 4)
    r = C.c_float(0)
    capi.check(_L().guda_sdot(n, x.ptr, incx, y.ptr, incy, C.byref(r), _sh(stream)), "Sdot")
    return np.float32(r.value)


def Ddot(n, x, incx, y, incy, stream=None):
    if incx == 0: raise BlasPanic("blas: zero x index increment")
    if incy == 0: raise BlasPanic("blas: zero y index increment")
    if n <= 0:
        if n == 0: return 0.0
        raise BlasPanic("blas: n < 0")
    _check_level1(n, x, incx, y, incy, 8)
    r = C.c_double(0)
    capi.check(_L().guda_ddot(n, x.ptr, incx, y.ptr, incy, C.byref(r), _sh(stream)), "Ddot")
    return r.value


def _check_dsdot(n, x, incx, y, incy):
    # compute/level1float32_dsdot.go:19-52 (order and strings)
    if incx == 0: raise BlasPanic("blas: zero x index increment")
    if incy == 0: raise BlasPanic("blas: zero y index increment")
    if n <= 0:
        if n == 0: return False
        raise BlasPanic("blas: n < 0")
    lx, ly = _len(x, 4), _len(y, 4)
    if incx == 1 and incy == 1:
        if lx < n: raise BlasPanic("blas: insufficient length of x")
        if ly < n: raise BlasPanic("blas: insufficient length of y")
        return True
    ix = (-n + 1) * incx if incx < 0 else 0
    iy = (-n + 1) * incy if incy < 0 else 0
    if ix >= lx or ix + (n - 1) * incx >= lx: raise BlasPanic("blas: insufficient length of x")
    if iy >= ly or iy + (n - 1) * incy >= ly: raise BlasPanic("blas: insufficient length of y")
    return True


def Dsdot(n, x, incx, y, incy, stream=None):       # compute/level1float32_dsdot.go:18
    if not _check_dsdot(n, x, incx, y, incy): return 0.0
    r = C.c_double(0)
    capi.check(_L().guda_dsdot(n, x.ptr, incx, y.ptr, incy, C.byref(r), _sh(stream)), "Dsdot")
    return r.value


def Sdsdot(n, alpha, x, incx, y, incy, stream=None):   # compute/level1float32_sdsdot.go:18
    if not _check_dsdot(n, x, incx, y, incy): return np.float32(0)
    r = C.c_float(0)
    capi.check(_L().guda_sdsdot(n, alpha, x.ptr, incx, y.ptr, incy, C.byref(r), _sh(stream)), "Sdsdot")
    return np.float32(r.value)


def Sscal(n, alpha, x, incx, stream=None):
    if incx < 1:
        if incx == 0: raise BlasPanic("blas: zero x index increment")
        return
    if n < 1:
        if n == 0: return
        raise BlasPanic("blas: n < 0")
    if (n - 1) * incx >= _len(x, 4): raise BlasPanic("blas: insufficient length of x")
    capi.check(_L().guda_sscal(n, alpha, x.ptr, incx, _sh(stream)), "Sscal")


def Dscal(n, alpha, x, incx, stream=None):
    if incx < 1:
        if incx == 0: raise BlasPanic("blas: zero x index increment")
        return
    if n < 1:
        if n == 0: return
        raise BlasPanic("blas: n < 0")
    if (n - 1) * incx >= _len(x, 8): raise BlasPanic("blas: insufficient length of x")
    capi.check(_L().guda_dscal(n, alpha, x.ptr, incx, _sh(stream)), "Dscal")


def Softmax(x, n, stream=None):                      # reduce.go:180-187
    x._slice_check(n)
    capi.check(_L().guda_softmax(x.ptr, n, _sh(stream)), "Softmax")


def LayerNorm(x, gamma, beta, n, eps, stream=None):  # fused.go:270-310
    x._slice_check(n)
    gp = gamma.ptr if (gamma is not None and gamma.ptr) else None
    bp = beta.ptr if (beta is not None and beta.ptr) else None
    capi.check(_L().guda_layer_norm(x.ptr, gp, bp, n, eps, _sh(stream)), "LayerNorm")


def Isamax(n, x, incx, stream=None):                # compute/level1float32.go:89-130
    if incx < 1:
        if incx == 0: raise BlasPanic("blas: zero x index increment")
        return -1
    if _len(x, 4) <= (n - 1) * incx: raise BlasPanic("blas: insufficient length of x")
    r = C.c_longlong(0)
    capi.check(_L().guda_isamax(n, x.ptr, incx, C.byref(r), _sh(stream)), "Isamax")
    return int(r.value)


def SgemmHost(transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream=None):
    """math.go:98 GEMM on HOST arrays (float32 numpy arrays, or integer addresses of pinned buffers): uploads pipelined under the
    GEMM by K chunks, C streamed back by row panels; blocks until c holds the result.  Same panics as Sgemm."""
    def addr(v):
        return v.ctypes.data if isinstance(v, np.ndarray) else int(v)
    if all(isinstance(v, np.ndarray) for v in (a, b, c)) and m > 0 and n > 0 and k >= 0 and min(lda, ldb, ldc) >= 1:
        la, lb, lc = a.size, b.size, c.size
        if (la < (k - 1) * lda + m) if transA else (la < (m - 1) * lda + k): raise BlasPanic("blas: insufficient length of a")
        if (lb < (n - 1) * ldb + k) if transB else (lb < (k - 1) * ldb + n): raise BlasPanic("blas: insufficient length of b")
        if lc < (m - 1) * ldc + n: raise BlasPanic("blas: insufficient length of c")
    capi.check(_L().guda_sgemm_host(int(bool(transA)), int(bool(transB)), m, n, k, alpha, addr(a), lda, addr(b), ldb, beta, addr(c), ldc,
                                    _sh(stream)), "GEMM")


# ---------------------------------------------------------------- Level 2
def Sgemv(trans, m, n, alpha, a, lda, x, incx, beta, y, incy, stream=None):    # compute/level2float32.go:244-308
    """Slice-length panics (which need len()) are raised here in the reference's order; the dimension / ld / increment panics
    come back from the C ABI with the same strings."""
    if m >= 0 and n >= 0 and lda >= max(1, n) and incx != 0 and incy != 0 and m > 0 and n > 0:
        lenx, leny = (m, n) if trans else (n, m)
        lx, ly = _len(x, 4), _len(y, 4)
        if (incx > 0 and (lenx - 1) * incx >= lx) or (incx < 0 and (1 - lenx) * incx >= lx): raise BlasPanic("blas: insufficient length of x")
        if (incy > 0 and (leny - 1) * incy >= ly) or (incy < 0 and (1 - leny) * incy >= ly): raise BlasPanic("blas: insufficient length of y")
        if _len(a, 4) < lda * (m - 1) + n: raise BlasPanic("blas: insufficient length of a")
    capi.check(_L().guda_sgemv(int(bool(trans)), m, n, alpha, a.ptr, lda, x.ptr, incx, beta, y.ptr, incy, _sh(stream)), "Sgemv")
    if trans and m > 0 and n > 0 and alpha != 0 and beta == 0 and incy == 1 and _len(y, 4) > n:
        # compute/asm/f32/gemv.go:61-64: `for i := range y { y[i] = 0 }` clears the whole slice, not only y[:n]
        capi.check(_L().guda_memset(y.ptr + 4 * n, 0, 4 * (_len(y, 4) - n), _sh(stream)), "Sgemv")


def Dgemv(trans, m, n, alpha, a, lda, x, incx, beta, y, incy, stream=None):    # compute/level2float64.go:232-300
    if m >= 0 and n >= 0 and lda >= max(1, n) and incx != 0 and incy != 0 and m > 0 and n > 0:
        lenx, leny = (m, n) if trans else (n, m)
        lx, ly = _len(x, 8), _len(y, 8)
        if (incx > 0 and (lenx - 1) * incx >= lx) or (incx < 0 and (1 - lenx) * incx >= lx): raise BlasPanic("blas: insufficient length of x")
        if (incy > 0 and (leny - 1) * incy >= ly) or (incy < 0 and (1 - leny) * incy >= ly): raise BlasPanic("blas: insufficient length of y")
        if _len(a, 8) < lda * (m - 1) + n: raise BlasPanic("blas: insufficient length of a")
    capi.check(_L().guda_dgemv(int(bool(trans)), m, n, alpha, a.ptr, lda, x.ptr, incx, beta, y.ptr, incy, _sh(stream)), "Dgemv")


def Dger(m, n, alpha, x, incx, y, incy, a, lda, stream=None):                  # compute/level2float64.go:24-65
    if m >= 0 and n >= 0 and lda >= max(1, n) and incx != 0 and incy != 0 and m > 0 and n > 0:
        lx, ly = _len(x, 8), _len(y, 8)
        if (incx > 0 and lx <= (m - 1) * incx) or (incx < 0 and lx <= (1 - m) * incx): raise BlasPanic("blas: insufficient length of x")
        if (incy > 0 and ly <= (n - 1) * incy) or (incy < 0 and ly <= (1 - n) * incy): raise BlasPanic("blas: insufficient length of y")
        if _len(a, 8) < lda * (m - 1) + n: raise BlasPanic("blas: insufficient length of a")
    capi.check(_L().guda_dger(m, n, alpha, x.ptr, incx, y.ptr, incy, a.ptr, lda, _sh(stream)), "Dger")


def Sger(m, n, alpha, x, incx, y, incy, a, lda, stream=None):                  # compute/level2float32.go:24-65
    if m >= 0 and n >= 0 and lda >= max(1, n) and incx != 0 and incy != 0 and m > 0 and n > 0:
        lx, ly = _len(x, 4), _len(y, 4)
        if (incx > 0 and lx <= (m - 1) * incx) or (incx < 0 and lx <= (1 - m) * incx): raise BlasPanic("blas: insufficient length of x")
        if (incy > 0 and ly <= (n - 1) * incy) or (incy < 0 and ly <= (1 - n) * incy): raise BlasPanic("blas: insufficient length of y")
        if _len(a, 4) < lda * (m - 1) + n: raise BlasPanic("blas: insufficient length of a")
    capi.check(_L().guda_sger(m, n, alpha, x.ptr, incx, y.ptr, incy, a.ptr, lda, _sh(stream)), "Sger")


def AXPY(alpha, x, y, n, stream=None):             # math.go:39-47
    Saxpy(n, alpha, x, 1, y, 1, stream)


def DOT(x, y, n, stream=None):                     # math.go:64-71
    return Sdot(n, x, 1, y, 1, stream)


def Scale(alpha, x, n, stream=None):               # math.go:170-177
    Sscal(n, alpha, x, 1, stream)


# ---------------------------------------------------------------- Level 3
def GEMM(transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream=None):   # math.go:98-129
    _dims_then_lengths(transA, transB, m, n, k, a, lda, b, ldb, c, ldc, 4)
    capi.check(_L().guda_sgemm(int(transA), int(transB), m, n, k, alpha, a.ptr, lda, b.ptr, ldb, beta, c.ptr, ldc, _sh(stream)), "GEMM")


def Dgemm(transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream=None):  # compute/dgemm.go:25
    _dims_then_lengths(transA, transB, m, n, k, a, lda, b, ldb, c, ldc, 8)
    capi.check(_L().guda_dgemm(int(transA), int(transB), m, n, k, alpha, a.ptr, lda, b.ptr, ldb, beta, c.ptr, ldc, _sh(stream)), "Dgemm")


def _dims_then_lengths(ta, tb, m, n, k, a, lda, b, ldb, c, ldc, item):
    # compute/sgemm.go:30-71 order: m, n, k, lda, ldb, ldc -- then (:74) quick return, then (:116-136) lengths
    if m < 0: raise BlasPanic("blas: m < 0")
    if n < 0: raise BlasPanic("blas: n < 0")
    if k < 0: raise BlasPanic("blas: k < 0")
    if lda < max(1, m if ta else k): raise BlasPanic("blas: bad leading dimension of A")
    if ldb < max(1, k if tb else n): raise BlasPanic("blas: bad leading dimension of B")
    if ldc < max(1, n): raise BlasPanic("blas: bad leading dimension of C")
    _check_gemm_lengths(ta, tb, m, n, k, a, lda, b, ldb, c, ldc, item)


Sgemm = GEMM


def FusedGEMMBiasReLU(transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, stream=None):   # fused.go:313
    if transA or transB or beta != 0:
        raise ErrNotSupported(capi.ERR_NOT_IMPLEMENTED, "FusedGEMMBiasReLU", "operation not supported")
    _dims_then_lengths(False, False, m, n, k, a, lda, b, ldb, c, ldc, 4)
    capi.check(_L().guda_sgemm_bias_relu(0, 0, m, n, k, alpha, a.ptr, lda, b.ptr, ldb, beta, c.ptr, ldc, bias.ptr, _sh(stream)),
               "FusedGEMMBiasReLU")


def FusedGEMMBiasGELU(transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, bias, stream=None):   # fused_gelu.go:65
    if transA or transB or beta != 0:
        raise ErrNotSupported(capi.ERR_NOT_IMPLEMENTED, "FusedGEMMBiasGELU", "operation not supported")
    _dims_then_lengths(False, False, m, n, k, a, lda, b, ldb, c, ldc, 4)
    capi.check(_L().guda_sgemm_bias_gelu(0, 0, m, n, k, alpha, a.ptr, lda, b.ptr, ldb, beta, c.ptr, ldc, bias.ptr, _sh(stream)),
               "FusedGEMMBiasGELU")


# ---------------------------------------------------------------- elementwise ops (math.go:133-169,230-272) and FusedKernel (fused.go)
def _ew_len(op, n, *ptrs):
    if n < 0 or any(p.size // 4 < n for p in ptrs):
        raise IndexError("%s: index out of range [%d]" % (op, n))       # the reference's slice indexing would panic


def Add(a, b, c, n, stream=None):                  # math.go:133
    _ew_len("Add", n, a, b, c)
    capi.check(_L().guda_add(a.ptr, b.ptr, c.ptr, n, _sh(stream)), "Add")


def Multiply(a, b, c, n, stream=None):             # math.go:152
    _ew_len("Multiply", n, a, b, c)
    capi.check(_L().guda_multiply(a.ptr, b.ptr, c.ptr, n, _sh(stream)), "Multiply")


def ReLU(x, n, stream=None):                       # math.go:236
    _ew_len("ReLU", n, x)
    capi.check(_L().guda_relu(x.ptr, n, _sh(stream)), "ReLU")


def Sigmoid(x, n, stream=None):                    # math.go:255 -> activations.go:9 SigmoidFloat32
    _ew_len("Sigmoid", n, x)
    capi.check(_L().guda_sigmoid(x.ptr, n, _sh(stream)), "Sigmoid")


def geluFloat32(x):                                # fused_gelu.go:44 -- the one function the reference hands to FusedKernel.Custom
    raise ErrNotSupported(capi.ERR_NOT_IMPLEMENTED, "geluFloat32", "device-side function: pass it to FusedKernel.Custom")


class FusedKernel:
    """fused.go:18-216: a list of ops applied to each element in one read-modify-write pass (guda_fused_execute)."""

    def __init__(self):
        self.ops = []                               # (type, alpha, broadcast, broadcast_dim)

    def Add(self, alpha):                           # fused.go:52
        self.ops.append((capi.FUSED_ADD, alpha, False, -1)); return self

    def AddBroadcast(self, alpha, broadcastDim):    # fused.go:63
        self.ops.append((capi.FUSED_ADD, alpha, True, broadcastDim)); return self

    def Multiply(self):                             # fused.go:74
        self.ops.append((capi.FUSED_MUL, 0.0, False, -1)); return self

    def AddScalar(self, alpha):                     # fused.go:82
        self.ops.append((capi.FUSED_ADD_SCALAR, alpha, False, -1)); return self

    def MulScalar(self, alpha):                     # fused.go:91
        self.ops.append((capi.FUSED_MUL_SCALAR, alpha, False, -1)); return self

    def ReLU(self):                                 # fused.go:100
        self.ops.append((capi.FUSED_RELU, 0.0, False, -1)); return self

    def Sigmoid(self):                              # FusedSigmoid, fused.go:40,200
        self.ops.append((capi.FUSED_SIGMOID, 0.0, False, -1)); return self

    def Custom(self, fn):                           # fused.go:108: host closures cannot run on the device; geluFloat32 can
        if fn is not geluFloat32:
            raise ErrNotSupported(capi.ERR_NOT_IMPLEMENTED, "FusedOp", "operation not supported")
        self.ops.append((capi.FUSED_GELU, 0.0, False, -1)); return self

    def Execute(self, x, others, n, *shapes, stream=None):      # fused.go:123
        others = list(others or [])
        _ew_len("FusedOp", n, x)
        ops = (capi.FusedOp * max(len(self.ops), 1))()
        oi = 0
        for k, (typ, alpha, bcast, dim) in enumerate(self.ops):
            bsize = bdim = 0
            if typ in (capi.FUSED_ADD, capi.FUSED_MUL) and oi < len(others):
                if bcast and dim >= 0 and oi < len(shapes):
                    shape = list(shapes[oi])
                    if dim >= len(shape):           # fused.go:155: the operand silently becomes 0 (Add) / 1 (Mul)
                        raise ErrNotSupported(capi.ERR_NOT_IMPLEMENTED, "FusedOp", "operation not supported")
                    bdim = int(shape[dim]); bsize = 1
                    for d_ in shape[dim:]:
                        bsize *= int(d_)
                    if bdim <= 0 or others[oi].size // 4 < bdim:
                        raise GUDAError(capi.ERR_INVALID_ARG, "FusedOp", "invalid shape")
                else:
                    _ew_len("FusedOp", n, others[oi])
                oi += 1
            ops[k] = capi.FusedOp(typ, alpha, bsize, bdim)
        ptrs = (C.c_void_p * max(len(others), 1))(*[o.ptr for o in others])
        capi.check(_L().guda_fused_execute(x.ptr, ptrs, len(others), n, ops, len(self.ops), _sh(stream)), "FusedOp")


def _reduce_axis(fn, op, x, shape, axis, out, stream):
    if len(shape) != 2:
        raise GUDAError(capi.ERR_INVALID_ARG, "FusedOp", "invalid shape")          # ErrInvalidShape, reduce.go:106
    m, n = int(shape[0]), int(shape[1])
    if axis not in (0, 1):
        raise GUDAError(capi.ERR_INVALID_ARG, "FusedOp", "invalid axis")           # ErrInvalidAxis, reduce.go:131
    _ew_len(op, m * n, x)
    _ew_len(op, n if axis == 0 else m, out)
    capi.check(getattr(_L(), fn)(x.ptr, m, n, axis, out.ptr, _sh(stream)), op)


def ReduceSum(x, shape, axis, out, stream=None):   # reduce.go:105
    _reduce_axis("guda_reduce_sum_axis", "ReduceSum", x, shape, axis, out, stream)


def ReduceMax(x, shape, axis, out, stream=None):   # reduce.go:139
    _reduce_axis("guda_reduce_max_axis", "ReduceMax", x, shape, axis, out, stream)


def TopK(x, n, k, stream=None):                    # reduce.go:247 -> (values, indices)
    _ew_len("TopK", n, x)
    kk = min(k, n)
    if kk <= 0:
        raise GUDAError(capi.ERR_INVALID_ARG, "FusedOp", "invalid parameter")      # ErrInvalidParameter
    vals = np.empty(kk, np.float32)
    idx = np.empty(kk, np.int64)
    capi.check(_L().guda_topk(x.ptr, n, k, vals.ctypes.data_as(C.POINTER(C.c_float)), idx.ctypes.data_as(C.POINTER(C.c_int64)),
                              _sh(stream)), "TopK")
    return vals, idx


def NewFusedKernel():                              # fused.go:45
    return FusedKernel()


def LinearReLU(x, alpha, beta, n, stream=None):    # fused.go:252: x = ReLU(alpha*x + beta)
    NewFusedKernel().MulScalar(alpha).AddScalar(beta).ReLU().Execute(x, None, n, stream=stream)


def GELU(x, n, stream=None):                       # fused.go:261
    NewFusedKernel().Custom(geluFloat32).Execute(x, None, n, stream=stream)


def AddBiasGELU(x, bias, n, stream=None):          # fused_gelu.go:12-41
    capi.check(_L().guda_add_bias_gelu(x.ptr, bias.ptr, n, bias.size // 4, _sh(stream)), "AddBiasGELU")


def AddBiasReLU(x, bias, n, stream=None):          # fused.go:220
    capi.check(_L().guda_add_bias_relu(x.ptr, bias.ptr, n, bias.size // 4, _sh(stream)), "AddBiasReLU")


def GEMMBFloat16(transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream=None):   # bfloat16.go:104
    capi.check(_L().guda_gemm_bf16(int(transA), int(transB), m, n, k, alpha, a.ptr, lda, b.ptr, ldb, beta, c.ptr, ldc, _sh(stream)),
               "GEMMBFloat16")


def GEMMFloat16(transA, transB, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc, stream=None):    # float16.go:201
    capi.check(_L().guda_gemm_f16(int(transA), int(transB), m, n, k, alpha, a.ptr, lda, b.ptr, ldb, beta, c.ptr, ldc, _sh(stream)),
               "GEMMFloat16")


def FusedGEMMBFloat16BiasGELU(m, n, k, alpha, a, lda, b, ldb, c, ldc, bias, stream=None):
    capi.check(_L().guda_gemm_bf16_bias_gelu(m, n, k, alpha, a.ptr, lda, b.ptr, ldb, c.ptr, ldc, bias.ptr, _sh(stream)),
               "FusedGEMMBFloat16BiasGELU")


def FusedGEMMBFloat16BiasReLU(m, n, k, alpha, a, lda, b, ldb, c, ldc, bias, stream=None):
    capi.check(_L().guda_gemm_bf16_bias_relu(m, n, k, alpha, a.ptr, lda, b.ptr, ldb, c.ptr, ldc, bias.ptr, _sh(stream)),
               "FusedGEMMBFloat16BiasReLU")


class ConvParams:
    """conv.go:9-31 ConvParams with the reference's field names."""
    FIELDS = ("BatchSize", "InChannels", "InHeight", "InWidth", "OutChannels", "KernelHeight", "KernelWidth",
              "StrideH", "StrideW", "PadH", "PadW", "DilationH", "DilationW", "UseBias")

    def __init__(self, **kw):
        for f in self.FIELDS:
            setattr(self, f, kw.get(f, 1 if f.startswith(("Stride", "Dilation")) else 0))

    def OutputHeight(self):                        # conv.go:54-57
        return (self.InHeight + 2 * self.PadH - ((self.KernelHeight - 1) * self.DilationH + 1)) // self.StrideH + 1

    def OutputWidth(self):                         # conv.go:60-63
        return (self.InWidth + 2 * self.PadW - ((self.KernelWidth - 1) * self.DilationW + 1)) // self.StrideW + 1

    def _c(self):
        return capi.ConvParams(*[int(getattr(self, f)) for f in self.FIELDS])


def Conv2D(input, kernel, bias, output, params, stream=None):          # conv.go:69
    """bias may be None / a zero DevicePtr (conv_test.go passes DevicePtr{})."""
    cp = params._c()
    bptr = bias.ptr if (bias is not None and bias.ptr) else None
    capi.check(_L().guda_conv2d(input.ptr, kernel.ptr, bptr, output.ptr, C.byref(cp), _sh(stream)), "Conv2D")


def Conv2DDirect(input, kernel, bias, output, params, stream=None):    # conv.go:231
    cp = params._c()
    bptr = bias.ptr if (bias is not None and bias.ptr) else None
    capi.check(_L().guda_conv2d_direct(input.ptr, kernel.ptr, bptr, output.ptr, C.byref(cp), _sh(stream)), "Conv2DDirect")


def SetConvRoute(route):
    """0 implicit GEMM (default), 1 materialised im2col, 2/3/4 implicit with the 64x128/64x64/128x128 tile forced,
    -1 default / environment (guda_conv_set_route)."""
    capi.check(_L().guda_conv_set_route(int(route)), "SetConvRoute")


def SetSgemmMode(mode):
    capi.check(_L().guda_set_sgemm_mode(mode), "SetSgemmMode")


def SetSgemmPipe(cfg=-1, lag=-1):
    """Exact-route kernel choice (bit-identical results either way): -1 automatic, 0 register-staged kernel only, 1.. a forced
    configuration of the TMA-fed pipeline (sgemm_pipe.cu; shapes that do not fit it still take the register-staged kernel)."""
    capi.check(_L().guda_set_sgemm_pipe(cfg, lag), "SetSgemmPipe")


def SetDgemmMode(mode):
    """0 = FP64 tensor pipe (DMMA) where eligible (default), 1 = SIMT DFMA kernel for every form."""
    capi.check(_L().guda_set_dgemm_mode(mode), "SetDgemmMode")


def SetGemm16Mode(mode):
    capi.check(_L().guda_set_gemm16_mode(mode), "SetGemm16Mode")


def FillUniform(ptr, rows, cols, ld, seed, row0=0, col0=0, full_cols=None, lo=-1.0, hi=1.0, dtype=np.float32, stream=None):
    full_cols = cols if full_cols is None else full_cols
    fn = _L().guda_fill_uniform_f32 if dtype == np.float32 else _L().guda_fill_uniform_f64
    capi.check(fn(ptr.ptr, rows, cols, ld, seed, row0, col0, full_cols, lo, hi, _sh(stream)), "FillUniform")


# ---- kernel-launch surface (guda.go:66-89,184-196,258-276,310-339; execution.go:9-86) ----
class Dim3:
    """guda.go:66-68."""

    def __init__(self, X=1, Y=1, Z=1):
        self.X, self.Y, self.Z = X, Y, Z

    def Size(self):                                   # guda.go:330-332
        return self.X * self.Y * self.Z


class ThreadID:
    """guda.go:73-78 with the index helpers of guda.go:310-327."""

    def __init__(self, BlockIdx, ThreadIdx, BlockDim, GridDim):
        self.BlockIdx, self.ThreadIdx, self.BlockDim, self.GridDim = BlockIdx, ThreadIdx, BlockDim, GridDim

    def Global(self): return self.BlockIdx.X * self.BlockDim.X + self.ThreadIdx.X
    def GlobalX(self): return self.Global()
    def GlobalY(self): return self.BlockIdx.Y * self.BlockDim.Y + self.ThreadIdx.Y
    def GlobalZ(self): return self.BlockIdx.Z * self.BlockDim.Z + self.ThreadIdx.Z


_LAUNCH_MSG = ("host closures cannot run on the device: use the device-kernel entry points (Add/Multiply/ReLU/Sigmoid, "
               "FusedKernel.Execute, LinearReLU, GELU, GEMMBFloat16/GEMMFloat16, the reductions)")


def _launch_refused(op, grid, block):
    if grid.Size() <= 0 or block.Size() <= 0:
        raise GUDAError(capi.ERR_INVALID_ARG, op, "grid and block dimensions must be positive")
    raise GUDAError(capi.ERR_NOT_IMPLEMENTED, op, _LAUNCH_MSG)


def Launch(kernel, grid, block, *args):               # guda.go:184-186: a Go closure per (block, thread) -- no device equivalent
    _launch_refused("Launch", grid, block)


def LaunchFunc(fn, grid, block, *args):               # guda.go:189-191
    _launch_refused("LaunchFunc", grid, block)


def LaunchStream(kernel, grid, block, stream, *args):     # guda.go:268-270
    _launch_refused("LaunchStream", grid, block)


def LaunchFuncStream(fn, grid, block, stream, *args):     # guda.go:273-275
    _launch_refused("LaunchFuncStream", grid, block)
