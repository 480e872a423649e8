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


# tolerance presets restated from the reference (tolerance.go:28-54, tolerance_arch.go:59-106,
# numerical_parity.go:139) -- amd64 values, which are the Base entries
TOL_DEFAULT = {"abs": 1e-7, "rel": 1e-5, "ulp": 4}
TOL_GEMM = {"abs": 1e-6, "rel": 1e-5, "ulp": 4}
TOL_REDUCE = {"abs": 1e-5, "rel": 1e-4, "ulp": 16}
TOL_GELU = {"abs": 1e-4, "rel": 1e-4, "ulp": 100}


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
    bound = 2.0 * max(k, 1) * u * (np.asarray(a_abs_b_abs, dtype=np.float64).ravel() +
                                   (0 if c_abs is None else np.asarray(c_abs, dtype=np.float64).ravel()))
    diff = np.abs(e.astype(np.float64) - g.astype(np.float64))
    near = np.array([O.near_equal(float(x), float(y), tol) for x, y in zip(e, g)]) if e.size <= 1 << 16 else \
        (diff <= np.maximum(tol["abs"], tol["rel"] * np.maximum(np.abs(e), np.abs(g))))
    bad = ~(near | (diff <= bound))
    v["n_outside_bound"] = int(bad.sum())
    return int(bad.sum()), v
