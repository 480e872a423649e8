"""Shared test helpers: golden loading, testblas tolerance, GUDA tolerance presets."""
import json
import math
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def d_tol_equal(a, b):
    """blas/testblas/common.go:22-38 dTolEqual (1e-14 normalised)."""
    if math.isnan(a) and math.isnan(b):
        return True
    if a == b:
        return True
    m = max(abs(a), abs(b))
    if m > 1:
        a /= m
        b /= m
    return abs(a - b) < 1e-14


def d_slice_tol_equal(a, b):
    return len(a) == len(b) and all(d_tol_equal(float(x), float(y)) for x, y in zip(a, b))


def flatten(rows):
    return np.array([v for r in rows for v in r], dtype=np.float64)


def transpose(rows):
    return [list(r) for r in zip(*rows)]


# Tolerances come from the product's tolerance framework (guda_b200/tolerance.py -> guda_b200/csrc/tolerance.cu; the
# restatement of tolerance.go:28-54, tolerance_arch.go:59-127 and numerical_parity.go:132-141 with the "sm_100a" entry), so the
# tests hold the GPU to the numbers a GUDA user would get from guda.GetOperationTolerance -- not to literals of their own.
# tests/test_tolerance.py pins those numbers on the reference's.  The dict form is what the oracle's checker takes.
def _tol_dict(t):
    return {"abs": float(t.AbsTol), "rel": float(t.RelTol), "ulp": int(t.ULPTol)}


def op_tolerance(op, arch=None):
    from guda_b200 import tolerance as T
    return T.GetOperationTolerance(op, arch)


def _std(op):
    from guda_b200 import tolerance as T
    s = T.StandardTolerances[op]
    return {"abs": s.AbsTol, "rel": s.RelTol, "ulp": s.ULPTol}


class _Lazy(dict):
    """dict filled from the framework on first use (importing tests/helpers.py must not need the built library)."""

    def __init__(self, make):
        super().__init__()
        self._make = make

    def _fill(self):
        if not len(self):
            self.update(self._make())

    def __getitem__(self, k):
        self._fill()
        return super().__getitem__(k)

    def keys(self):
        self._fill()
        return super().keys()


def _default():
    from guda_b200 import tolerance as T
    return _tol_dict(T.DefaultTolerance())


TOL_DEFAULT = _Lazy(_default)                                          # tolerance.go:28-36
TOL_GEMM = _Lazy(lambda: _tol_dict(op_tolerance("gemm")))              # tolerance_arch.go:59-66 (+ sm_100a: same three numbers)
TOL_REDUCE = _Lazy(lambda: _tol_dict(op_tolerance("reduce_sum")))      # tolerance_arch.go:99-106
TOL_GELU = _Lazy(lambda: _std("gelu"))                                 # numerical_parity.go:139


def gemm_parity(expected, actual, tol, a_abs_b_abs=None, k=0, u=2.0 ** -24, c_abs=None):
    """SURVEY.md section 8(d) rule: an element passes if the reference's Float32NearEqual passes
    OR |delta| <= 2*K*u*(sum_l |a_il||b_lj| + |beta*c_ij|) (the classical forward-error bound at depth K).
    Returns (n_fail, stats).  a_abs_b_abs is |alpha| * |A| @ |B| computed in float64 by the caller."""
    from oracle import oracle as O
    e = np.asarray(expected, dtype=np.float32).ravel()
    g = np.asarray(actual, dtype=np.float32).ravel()
    v = O.verify_f32(e, g, tol)
    if v["num_errors"] == 0 or a_abs_b_abs is None:
        return v["num_errors"], v
    from guda_b200 import tolerance as T
    bound = T.k_scaled_bound(op_tolerance("gemm" if u > 2.0 ** -40 else "dgemm"), k, np.asarray(a_abs_b_abs, dtype=np.float64).ravel(), u,
                             None if c_abs is None else np.asarray(c_abs, dtype=np.float64).ravel())
    diff = np.abs(e.astype(np.float64) - g.astype(np.float64))
    near = np.array([O.near_equal(float(x), float(y), tol) for x, y in zip(e, g)]) if e.size <= 1 << 16 else \
        (diff <= np.maximum(tol["abs"], tol["rel"] * np.maximum(np.abs(e), np.abs(g))))
    bad = ~(near | (diff <= bound))
    v["n_outside_bound"] = int(bad.sum())
    return int(bad.sum()), v
