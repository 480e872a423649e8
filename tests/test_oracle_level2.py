"""The level-2 oracle (oracle/level2.c) against the reference's own golden vectors:
compute/asm/f32/gemv_test.go (SgemvCases, harness :508-615) and compute/asm/f32/ge_test.go (gerTests, harness :146-219),
with the reference's tolerance (1e-6 absolute-or-relative, util_test.go:27-30) and its guard-band / read-only checks."""
import itertools
import json
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "level2_cases.json")))
TOL = 1e-6
GD = 4


def same_approx(got, want, tol=TOL):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    d = np.abs(got - want)
    both_nan = np.isnan(got) & np.isnan(want)                       # scalar.Same: NaN matches NaN
    return bool(np.all(both_nan | (got == want) | (d <= tol) | (d <= tol * np.maximum(np.abs(got), np.abs(want)))))


def guard_inc(vec, gdval, inc, gdlen=GD):
    inc = abs(inc)
    g = np.full(len(vec) * inc + 2 * gdlen, gdval, np.float32)
    g[gdlen:gdlen + len(vec) * inc:inc] = np.asarray(vec, np.float32)
    return g


def guards_intact(g, vec_len, gdval, inc, gdlen=GD):
    inc = abs(inc)
    mask = np.ones(g.size, bool)
    mask[gdlen:gdlen + vec_len * inc:inc] = False
    return bool(np.all(g[mask] == np.float32(gdval)))


GEMV_IDS = ["%dx%d" % (c["m"], c["n"]) for c in GOLD["gemv"]]


@pytest.mark.parametrize("case", GOLD["gemv"], ids=GEMV_IDS)
@pytest.mark.parametrize("trans", [False, True])
def test_sgemv_golden_unit_and_strided(case, trans):
    m, n = case["m"], case["n"]
    x0, y0 = (case["y"], case["x"]) if trans else (case["x"], case["y"])
    a = np.asarray(case["A"], np.float32)
    for sub in case["Trans" if trans else "NoTrans"]:
        for incx, incy in itertools.product([1, -1, 2, 3, 90], repeat=2):
            want = sub["want"]
            if incx < 0 and incy < 0:
                want = sub["wantRevXY"]
            elif incx < 0:
                want = sub["wantRevX"]
            elif incy < 0:
                want = sub["wantRevY"]
            xg, yg = guard_inc(x0, 0.5, incx), guard_inc(y0, 1.5, incy)
            x, y = xg[GD:xg.size - GD], yg[GD:yg.size - GD]
            y2 = y.copy()
            O.f32_gemv(trans, m, n, sub["alpha"], a, n, x, incx, sub["beta"], y, incy)          # the level the goldens are stated at
            got = y[::abs(incy)][:len(want)]
            assert same_approx(got, want), (m, n, trans, sub["alpha"], sub["beta"], incx, incy, got, want)
            # Sgemv (level2float32.go:244-308) differs only by its alpha == 0 shortcut: 0 * NaN never happens there
            O.sgemv(trans, m, n, sub["alpha"], a, n, x, incx, sub["beta"], y2, incy)
            if sub["alpha"] != 0:
                assert np.array_equal(y2, y, equal_nan=True)
            else:
                y0f = np.asarray(y0, np.float32)
                exp = y0f if sub["beta"] == 1 else (np.zeros_like(y0f) if sub["beta"] == 0 else y0f * np.float32(sub["beta"]))
                assert np.array_equal(y2[::abs(incy)][:len(y0)], exp)
            assert guards_intact(xg, len(x0), 0.5, incx) and guards_intact(yg, len(y0), 1.5, incy)
            assert np.array_equal(x[::abs(incx)], np.asarray(x0, np.float32))
    assert np.array_equal(a, np.asarray(case["A"], np.float32), equal_nan=True)


@pytest.mark.parametrize("case", GOLD["ger"], ids=["%dx%d" % (len(c["x"]), len(c["y"])) for c in GOLD["ger"]])
def test_sger_golden(case):
    m, n = len(case["x"]), len(case["y"])
    x0, y0, a0 = (np.asarray(case[k], np.float32) for k in ("x", "y", "a"))
    for ax, ay in itertools.product(range(2), repeat=2):              # align2: shifted starts
        xb = np.concatenate([np.zeros(ax, np.float32), x0]); yb = np.concatenate([np.zeros(ay, np.float32), y0])
        a = a0.copy()
        O.sger(m, n, 1.0, xb[ax:], 1, yb[ay:], 1, a, n)
        assert same_approx(a, case["want"]), (m, n, a, case["want"])
    for incx, incy in itertools.product([1, 2], repeat=2):
        xg, yg = guard_inc(x0, -0.5, incx), guard_inc(y0, 1.5, incy)
        a = a0.copy()
        O.sger(m, n, 3.5, xg[GD:xg.size - GD], incx, yg[GD:yg.size - GD], incy, a, n)
        want = (np.float32(3.5) * x0[:, None] * y0[None, :]).astype(np.float32).ravel() + a0     # ge_test.go:199
        assert same_approx(a, want)
        assert guards_intact(xg, m, -0.5, incx) and guards_intact(yg, n, 1.5, incy)


def _ref_gemv(trans, m, n, alpha, a, lda, x, incx, beta, y, incy):
    """float64 model of y = alpha*op(A)x + beta*y with BLAS increment semantics (negative = reversed traversal)."""
    A = np.asarray(a, np.float64)[: lda * (m - 1) + n]
    A = np.array([A[i * lda:i * lda + n] for i in range(m)])
    lx, ly = (m, n) if trans else (n, m)
    xs = np.asarray(x, np.float64)[::abs(incx)][:lx]
    ys = np.asarray(y, np.float64)[::abs(incy)][:ly]
    if incx < 0:
        xs = xs[::-1]
    if incy < 0:
        ys = ys[::-1]
    r = alpha * ((A.T if trans else A) @ xs) + (beta * ys if beta != 0 else 0.0)
    return r[::-1] if incy < 0 else r


@pytest.mark.parametrize("trans", [False, True])
@pytest.mark.parametrize("m,n", [(1, 1), (3, 70), (70, 3), (33, 129), (64, 64), (5, 31), (5, 32), (5, 63)])
def test_sgemv_random_against_float64(trans, m, n):
    lda = n + 3
    a = O.uniform_f32(m * lda, 7); lx, ly = (m, n) if trans else (n, m)
    for incx, incy, alpha, beta in [(1, 1, 1.0, 0.0), (1, 1, -0.5, 2.0), (2, 1, 1.5, 1.0), (1, -3, 2.0, 0.5), (-2, 2, 1.0, 0.0)]:
        x = O.uniform_f32(lx * abs(incx), 8); y = O.uniform_f32(ly * abs(incy), 9)
        want = _ref_gemv(trans, m, n, alpha, a, lda, x, incx, beta, y, incy)
        got = O.sgemv(trans, m, n, alpha, a, lda, x, incx, beta, y.copy(), incy)[::abs(incy)][:ly]
        assert np.allclose(got, want, rtol=1e-5, atol=1e-5 * max(m, n))


def test_sgemv_quick_returns_and_panics():
    a = np.ones(6, np.float32); x = np.ones(3, np.float32); y = np.full(2, 7.0, np.float32)
    for args, msg in [((False, -1, 3, 1.0, a, 3, x, 1, 0.0, y, 1), "blas: m < 0"), ((False, 2, -1, 1.0, a, 3, x, 1, 0.0, y, 1), "blas: n < 0"),
                      ((False, 2, 3, 1.0, a, 2, x, 1, 0.0, y, 1), "blas: bad leading dimension of A"),
                      ((False, 2, 3, 1.0, a, 3, x, 0, 0.0, y, 1), "blas: zero x index increment"),
                      ((False, 2, 3, 1.0, a, 3, x, 1, 0.0, y, 0), "blas: zero y index increment"),
                      ((False, 2, 3, 1.0, a, 3, x[:2], 1, 0.0, y, 1), "blas: insufficient length of x"),
                      ((False, 2, 3, 1.0, a, 3, x, 1, 0.0, y[:1], 1), "blas: insufficient length of y"),
                      ((False, 2, 3, 1.0, a[:5], 3, x, 1, 0.0, y, 1), "blas: insufficient length of a")]:
        with pytest.raises(O.OraclePanic, match=msg.replace("(", r"\(")):
            O.sgemv(*args)
    assert np.array_equal(O.sgemv(False, 2, 3, 0.0, a, 3, x, 1, 1.0, y.copy(), 1), y)          # alpha 0, beta 1: untouched
    assert np.array_equal(O.sgemv(False, 2, 3, 0.0, a, 3, x, 1, 0.5, y.copy(), 1), y * 0.5)    # alpha 0: Sscal(beta)
    nan = np.array([np.nan, np.inf], np.float32)
    assert np.array_equal(O.sgemv(False, 2, 3, 0.0, a, 3, x, 1, 0.0, nan.copy(), 1), np.zeros(2, np.float32))
    assert np.array_equal(O.sgemv(False, 2, 3, 1.0, a, 3, x, 1, 0.0, nan.copy(), 1), np.full(2, 3.0, np.float32))   # beta 0 ignores y
    # GemvT with beta == 0 and unit incY clears the whole y slice, beyond n (gemv.go:61-64)
    ylong = np.full(5, 9.0, np.float32)
    O.sgemv(True, 2, 3, 1.0, a, 3, np.ones(2, np.float32), 1, 0.0, ylong, 1)
    assert np.array_equal(ylong, np.array([2, 2, 2, 0, 0], np.float32))


def test_sger_quick_returns_and_panics():
    a = np.ones(6, np.float32); x = np.ones(2, np.float32); y = np.ones(3, np.float32)
    for args, msg in [((-1, 3, 1.0, x, 1, y, 1, a, 3), "blas: m < 0"), ((2, 3, 1.0, x, 1, y, 1, a, 2), "blas: bad leading dimension of A"),
                      ((2, 3, 1.0, x, 0, y, 1, a, 3), "blas: zero x index increment"), ((2, 3, 1.0, x[:1], 1, y, 1, a, 3), "blas: insufficient length of x"),
                      ((2, 3, 1.0, x, 1, y, 2, a, 3), "blas: insufficient length of y"), ((2, 3, 1.0, x, 1, y, 1, a[:5], 3), "blas: insufficient length of a")]:
        with pytest.raises(O.OraclePanic, match=msg):
            O.sger(*args)
    assert np.array_equal(O.sger(2, 3, 0.0, x, 1, y, 1, a.copy(), 3), a)
    # unfused: (alpha*x)*y rounded, then added
    xs = np.array([1.0000001], np.float32); ys = np.array([1.0000001], np.float32); aa = np.array([-1.0], np.float32)
    got = O.sger(1, 1, 1.0, xs, 1, ys, 1, aa.copy(), 1)[0]
    assert got == np.float32(np.float32(xs[0] * ys[0]) + aa[0])


# ---------------------------------------------------------------- fp64
DGEMV = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "dgemv_cases.json")))["cases"]


@pytest.mark.parametrize("case", DGEMV, ids=[c["Name"] for c in DGEMV])
def test_dgemv_testblas_cases(case):
    """blas/testblas/dgemv.go: DgemvCases with the harness's increment flips (dgemvcomp :643-683), dSliceTolEqual (1e-14 normalised)."""
    trans = "NoTrans" not in case["tA"]
    a = np.array(case["A"], np.float64).ravel()
    for sub in case["Subcases"]:
        incx = -case["incX"] if sub.get("mulXNeg1") else case["incX"]
        incy = -case["incY"] if sub.get("mulYNeg1") else case["incY"]
        x = np.array(case["x"], np.float64); y = np.array(case["y"], np.float64)
        O.dgemv(trans, case["m"], case["n"], sub["alpha"], a, case["n"], x, incx, sub["beta"], y, incy)
        ans = np.array(sub["ans"], np.float64)
        assert np.all(np.abs(y - ans) <= 1e-14 * np.maximum(1.0, np.maximum(np.abs(y), np.abs(ans)))), (case["Name"], sub)
        assert np.array_equal(x, np.array(case["x"], np.float64))


def test_dger_is_the_unfused_elementwise_form():
    """compute/asm/f64/ger_amd64.s: fl(fl(alpha*x_i)*y_j) + a_ij -- numpy float64 arithmetic is the exact model."""
    rng = np.random.default_rng(3)
    for m, n, incx, incy in [(1, 1, 1, 1), (5, 3, 1, 1), (7, 9, 2, 3), (6, 4, -1, 1), (4, 6, 1, -2)]:
        x = rng.standard_normal((m - 1) * abs(incx) + 1); y = rng.standard_normal((n - 1) * abs(incy) + 1); a = rng.standard_normal(m * (n + 1))
        xs, ys = x[::abs(incx)][:m], y[::abs(incy)][:n]
        if incx < 0: xs = xs[::-1]
        if incy < 0: ys = ys[::-1]
        want = a.copy().reshape(m, n + 1)
        want[:, :n] = ((1.7 * xs)[:, None] * ys[None, :]) + want[:, :n]
        got = O.dger(m, n, 1.7, x, incx, y, incy, a.copy(), n + 1)
        assert np.array_equal(got, want.ravel())
    with pytest.raises(O.OraclePanic, match="blas: insufficient length of a"):
        O.dger(2, 3, 1.0, np.ones(2), 1, np.ones(3), 1, np.ones(5), 3)


def test_dgemv_t_is_the_sequential_unfused_chain():
    rng = np.random.default_rng(4)
    m, n = 37, 21
    a = rng.standard_normal(m * n); x = rng.standard_normal(m); y = rng.standard_normal(n)
    want = y * 0.5
    for i in range(m):
        want = want + (1.3 * x[i]) * a[i * n:(i + 1) * n]                 # numpy: one rounding per operation, row order
    assert np.array_equal(O.dgemv(True, m, n, 1.3, a, n, x, 1, 0.5, y.copy(), 1), want)
