"""GUDA's architecture-aware tolerance framework, host mirror (same names as the reference's package-level API).

  tolerance.go:10-191        ToleranceConfig, Default/Strict/RelaxedTolerance, Float32NearEqual, Float32ULPDiff,
                             VerifyFloat32Array -> VerificationResult{IsAcceptable, String}
  tolerance_arch.go:8-127    ArchToleranceConfig, GetArchTolerance, mergeTolerances, GEMM/Conv/ReduceArchTolerance,
                             GetOperationTolerance, IsARM64
  numerical_parity.go:103-141  OperationTolerance, StandardTolerances

The numbers live in ONE place, the C library (guda_b200/csrc/tolerance.cu); this module reads them through the C ABI, so the Go
shim, the C++ header and these Python names cannot drift apart.  The reference keys overrides on runtime.GOARCH; results that
come out of this library were computed on an sm_100a device, so ARCH defaults to "sm_100a" -- the new architecture entry,
which carries the K-scaled GEMM bound, the 16-bit-ULP bound and the fp64 bounds the reference has no field for.
VerifyFloat32Array takes host arrays (numpy) or two DevicePtr (then the comparison is one streaming pass on the device)."""
import ctypes as C
import dataclasses

import numpy as np

from . import capi

ARCH = "sm_100a"
MAX_INT32 = 0x7FFFFFFF


@dataclasses.dataclass
class ToleranceConfig:                                  # tolerance.go:10-26 (+ the sm_100a fields, zero = not applicable)
    AbsTol: float = 0.0
    RelTol: float = 0.0
    ULPTol: int = 0
    CheckNaN: bool = False
    CheckInf: bool = False
    KScale: float = 0.0          # |delta| <= KScale * K * u * sum_l |a_il||b_lj| also passes
    HalfULPTol: int = 0          # bf16 / fp16 outputs: ULPs of the 16-bit format
    Abs64: float = 0.0           # fp64 results
    Rel64: float = 0.0

    def _c(self):
        return capi.Tolerance(np.float32(self.AbsTol), np.float32(self.RelTol), int(self.ULPTol), int(self.CheckNaN), int(self.CheckInf),
                              np.float32(self.KScale), int(self.HalfULPTol), float(self.Abs64), float(self.Rel64))

    @staticmethod
    def _from(t):
        return ToleranceConfig(float(t.abs_tol), float(t.rel_tol), int(t.ulp_tol), bool(t.check_nan), bool(t.check_inf), float(t.k_scale),
                               int(t.half_ulp_tol), float(t.abs64), float(t.rel64))


def _preset(which):
    t = capi.Tolerance()
    capi.check(capi.load().guda_tolerance_preset(which, C.byref(t)), "TolerancePreset")
    return ToleranceConfig._from(t)


def DefaultTolerance(): return _preset(0)               # tolerance.go:28-36
def StrictTolerance(): return _preset(1)                # tolerance.go:39-45
def RelaxedTolerance(): return _preset(2)               # tolerance.go:48-54


def Float32ULPDiff(a, b):                               # tolerance.go:100-118
    ab, bb = int(np.float32(a).view(np.uint32)), int(np.float32(b).view(np.uint32))
    if (ab ^ bb) & 0x80000000:
        return MAX_INT32
    return abs(ab - bb)


def Float32NearEqual(a, b, tol):                        # tolerance.go:57-97
    a, b = np.float32(a), np.float32(b)
    if tol.CheckNaN and np.isnan(a) and np.isnan(b):
        return True
    if tol.CheckInf and np.isinf(a) and np.isinf(b) and (a > 0) == (b > 0):
        return True
    if a == b:
        return True
    with np.errstate(all="ignore"):
        diff = abs(float(np.float32(a - b)))
    if diff <= float(np.float32(tol.AbsTol)):
        return True
    larger = max(abs(float(a)), abs(float(b)))
    if diff <= larger * float(np.float32(tol.RelTol)):
        return True
    return bool(tol.ULPTol > 0 and Float32ULPDiff(a, b) <= tol.ULPTol)


@dataclasses.dataclass
class VerificationResult:                               # tolerance.go:121-128
    MaxAbsError: float = 0.0
    MaxRelError: float = 0.0
    MaxULPError: int = 0
    NumErrors: int = 0
    TotalItems: int = 0
    FirstError: int = -1

    def IsAcceptable(self, tol):                        # tolerance.go:177-182
        return bool(self.NumErrors == 0 or (self.MaxAbsError <= np.float32(tol.AbsTol) and self.MaxRelError <= np.float32(tol.RelTol)
                                            and self.MaxULPError <= tol.ULPTol))

    def String(self):                                   # tolerance.go:185-200
        if self.NumErrors == 0:
            return "PASS: All values match within tolerance"
        rate = 100.0 * self.NumErrors / max(self.TotalItems, 1)
        return ("FAIL: %d/%d values differ (%.2f%%)\n  Max absolute error: %e\n  Max relative error: %e\n  Max ULP error: %d\n"
                "  First error at index: %d" % (self.NumErrors, self.TotalItems, rate, self.MaxAbsError, self.MaxRelError,
                                                self.MaxULPError, self.FirstError))


def VerifyFloat32Array(expected, actual, tol, n=None, stream=None):     # tolerance.go:131-174
    """Host arrays: the reference's loop.  Two DevicePtr (n = element count, default the smaller buffer): guda_verify_f32."""
    if hasattr(expected, "ptr") and hasattr(actual, "ptr"):
        n = min(expected.size, actual.size) // 4 if n is None else n
        r = capi.VerifyResult()
        capi.check(capi.load().guda_verify_f32(expected.ptr, actual.ptr, n, C.byref(tol._c()), C.byref(r),
                                               None if stream is None else stream.handle), "VerifyFloat32Array")
        return VerificationResult(float(r.max_abs_error), float(r.max_rel_error), int(r.max_ulp_error), int(r.num_errors),
                                  int(r.total_items), int(r.first_error))
    e = np.asarray(expected, dtype=np.float32).ravel()
    a = np.asarray(actual, dtype=np.float32).ravel()
    res = VerificationResult(TotalItems=e.size)
    if e.size != a.size:
        res.NumErrors = e.size                          # tolerance.go:137-141
        return res
    for i in range(e.size):
        if Float32NearEqual(e[i], a[i], tol):
            continue
        res.NumErrors += 1
        if res.FirstError == -1:
            res.FirstError = i
        with np.errstate(all="ignore"):
            ad = np.float32(abs(float(np.float32(e[i] - a[i]))))
            if ad > res.MaxAbsError:
                res.MaxAbsError = float(ad)
            if e[i] != 0:
                rd = np.float32(ad / np.float32(abs(float(e[i]))))
                if rd > res.MaxRelError:
                    res.MaxRelError = float(rd)
        ud = Float32ULPDiff(e[i], a[i])
        if ud > res.MaxULPError:
            res.MaxULPError = ud
    return res


def mergeTolerances(base, override):                    # tolerance_arch.go:41-57
    out = capi.Tolerance()
    capi.check(capi.load().guda_merge_tolerances(C.byref(base._c()), C.byref(override._c()), C.byref(out)), "mergeTolerances")
    return ToleranceConfig._from(out)


@dataclasses.dataclass
class ArchToleranceConfig:                              # tolerance_arch.go:8-17 + the SM100a override
    Base: ToleranceConfig
    AMD64: ToleranceConfig = None
    ARM64: ToleranceConfig = None
    Generic: ToleranceConfig = None
    SM100a: ToleranceConfig = None


def GetArchTolerance(config, arch=None):                # tolerance_arch.go:20-38
    arch = ARCH if arch is None else arch
    over = {"sm_100a": config.SM100a, "amd64": config.AMD64, "arm64": config.ARM64, "arm64be": config.ARM64}.get(arch, config.Generic)
    return config.Base if over is None else mergeTolerances(config.Base, over)


def GetOperationTolerance(operation, arch=None):        # tolerance_arch.go:116-127
    t = capi.Tolerance()
    capi.check(capi.load().guda_get_operation_tolerance(operation.encode(), None if arch is None else arch.encode(), C.byref(t)),
               "GetOperationTolerance")
    return ToleranceConfig._from(t)


def _arch_config(op):
    base = GetOperationTolerance(op, "amd64")
    def over(arch):
        t = GetOperationTolerance(op, arch)
        return None if t == base else t
    return ArchToleranceConfig(Base=base, ARM64=over("arm64"), Generic=over("generic"), SM100a=over("sm_100a"))


def __getattr__(name):                                  # GEMMArchTolerance / ConvArchTolerance / ReduceArchTolerance (tolerance_arch.go:59-113)
    ops = {"GEMMArchTolerance": "gemm", "ConvArchTolerance": "conv2d", "ReduceArchTolerance": "reduce_sum"}
    if name in ops:
        return _arch_config(ops[name])
    if name == "StandardTolerances":
        return {op: _standard(op) for op in ("add", "multiply", "gemm", "conv2d", "reduce_sum", "softmax", "gelu", "exp")}
    raise AttributeError(name)


def IsARM64():                                          # tolerance_arch.go:130-132: the compute device is never ARM64
    return False


@dataclasses.dataclass
class OperationTolerance:                               # numerical_parity.go:124-129
    Name: str
    AbsTol: float
    RelTol: float
    ULPTol: int


def _standard(op):
    a, r, u = C.c_float(0), C.c_float(0), C.c_int(0)
    capi.check(capi.load().guda_standard_tolerance(op.encode(), C.byref(a), C.byref(r), C.byref(u)), "StandardTolerances")
    return OperationTolerance(op, float(a.value), float(r.value), int(u.value))


# ---- helpers the parity tests share (SURVEY.md 8d "Parity procedure at scale") ----
def k_scaled_bound(tol, k, abs_a_abs_b, unit_roundoff=2.0 ** -24, abs_beta_c=None):
    """tol.KScale * K * u * (sum_l |a_il||b_lj| [+ |beta * c_ij|]) as a float64 array (0 where the entry has no K-scaled bound)."""
    s = np.asarray(abs_a_abs_b, dtype=np.float64)
    if abs_beta_c is not None:
        s = s + np.asarray(abs_beta_c, dtype=np.float64)
    return float(tol.KScale) * max(int(k), 1) * unit_roundoff * s


def half_ulp_bound(tol, want, mant_bits, min_exp):
    """tol.HalfULPTol ULPs of a 16-bit format with `mant_bits` explicit mantissa bits (bf16: 7, fp16: 10) around `want`."""
    w = np.maximum(np.abs(np.asarray(want, dtype=np.float64)), 2.0 ** min_exp)
    return tol.HalfULPTol * w * 2.0 ** -mant_bits
