"""GPU parity for SGEMV / SGER (guda_b200/csrc/level2.cu) through the C ABI: bit-exact against the oracle's restatement of
compute/asm/f32/gemv.go + dotunitary_avx2_amd64.s / dotinc_amd64.s / axpyunitaryto_amd64.s and ge_amd64.s, for every
increment sign, plus the reference's golden vectors (gemv_test.go, ge_test.go) under its own 1e-6 tolerance."""
import itertools
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "level2_cases.json")))
GD = 4


@pytest.fixture(scope="module")
def g():
    import guda_b200 as G
    G.Init(0)
    return G


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


def same_approx(got, want, tol=1e-6):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    d = np.abs(got - want)
    return bool(np.all((np.isnan(got) & np.isnan(want)) | (got == want) | (d <= tol) | (d <= tol * np.maximum(np.abs(got), np.abs(want)))))


def guard_inc(vec, gdval, inc):
    inc = abs(inc)
    gd = np.full(len(vec) * inc + 2 * GD, gdval, np.float32)
    gd[GD:GD + len(vec) * inc:inc] = np.asarray(vec, np.float32)
    return gd


def gpu_gemv(g, trans, m, n, alpha, a, lda, x, incx, beta, y, incy):
    da, dx, dy = g.to_device(a), g.to_device(x), g.to_device(y)
    g.Sgemv(trans, m, n, alpha, da, lda, dx, incx, beta, dy, incy)
    g.Synchronize()
    out = dy.Float32()
    assert np.array_equal(da.Float32(), a, equal_nan=True) and np.array_equal(dx.Float32(), x, equal_nan=True)   # inputs read-only
    g.Free(da); g.Free(dx); g.Free(dy)
    return out


def gpu_ger(g, m, n, alpha, x, incx, y, incy, a, lda):
    da, dx, dy = g.to_device(a), g.to_device(x), g.to_device(y)
    g.Sger(m, n, alpha, dx, incx, dy, incy, da, lda)
    g.Synchronize()
    out = da.Float32()
    assert np.array_equal(dx.Float32(), x, equal_nan=True) and np.array_equal(dy.Float32(), y, equal_nan=True)
    g.Free(da); g.Free(dx); g.Free(dy)
    return out


@pytest.mark.parametrize("case", GOLD["gemv"], ids=["%dx%d" % (c["m"], c["n"]) for c in GOLD["gemv"]])
@pytest.mark.parametrize("trans", [False, True])
def test_sgemv_golden_vectors(g, O, case, trans):
    """gemv_test.go's harness (guards, increments -1/1/2/3/90).  The goldens are stated for package f32's GemvN/GemvT; Sgemv adds
    the alpha == 0 shortcut (level2float32.go:286-299), so the NaN-in-A rows of alpha == 0 subcases are checked via the oracle."""
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
            x, y = xg[GD:xg.size - GD].copy(), yg[GD:yg.size - GD].copy()
            got = gpu_gemv(g, trans, m, n, sub["alpha"], a, n, x, incx, sub["beta"], y, incy)
            ref = O.sgemv(trans, m, n, sub["alpha"], a, n, x, incx, sub["beta"], y.copy(), incy)
            assert np.array_equal(got, ref, equal_nan=True), (m, n, trans, sub["alpha"], sub["beta"], incx, incy)   # guards included
            if sub["alpha"] != 0:
                assert same_approx(got[::abs(incy)][:len(want)], want)


@pytest.mark.parametrize("case", GOLD["ger"], ids=["%dx%d" % (len(c["x"]), len(c["y"])) for c in GOLD["ger"]])
def test_sger_golden_vectors(g, O, case):
    m, n = len(case["x"]), len(case["y"])
    x0, y0, a0 = (np.asarray(case[k], np.float32) for k in ("x", "y", "a"))
    got = gpu_ger(g, m, n, 1.0, x0, 1, y0, 1, a0, n)
    assert same_approx(got, case["want"]) and np.array_equal(got, O.sger(m, n, 1.0, x0, 1, y0, 1, a0.copy(), n), equal_nan=True)
    for incx, incy in itertools.product([1, 2, -1, -3], repeat=2):
        xg, yg = guard_inc(x0, -0.5, incx), guard_inc(y0, 1.5, incy)
        x, y = xg[GD:xg.size - GD].copy(), yg[GD:yg.size - GD].copy()
        got = gpu_ger(g, m, n, 3.5, x, incx, y, incy, a0, n)
        assert np.array_equal(got, O.sger(m, n, 3.5, x, incx, y, incy, a0.copy(), n), equal_nan=True)
        if incx > 0 and incy > 0:
            want = (np.float32(3.5) * x0[:, None] * y0[None, :]).astype(np.float32).ravel() + a0     # ge_test.go:199
            assert same_approx(got, want)


# every tail branch of the AVX2 dot: 32-blocked body, 16/8/4-wide tails, scalar tail
N_SHAPES = [(1, 1), (3, 5), (7, 31), (5, 32), (9, 33), (4, 47), (6, 60), (13, 63), (2, 64), (17, 100), (33, 255), (64, 256), (130, 1000),
            (257, 4099), (1025, 517), (3, 20000)]


@pytest.mark.parametrize("m,n", N_SHAPES)
@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (0.75, -1.5), (1.0, 1.0)])
def test_sgemv_n_unit_bit_exact(g, O, m, n, alpha, beta):
    lda = n + (m % 3)
    a = O.uniform_f32(m * lda, 3); x = O.uniform_f32(n, 4); y = O.uniform_f32(m, 5)
    want = O.sgemv(False, m, n, alpha, a, lda, x, 1, beta, y.copy(), 1)
    got = gpu_gemv(g, False, m, n, alpha, a, lda, x, 1, beta, y, 1)
    assert np.array_equal(got, want), int(np.sum(got != want))


@pytest.mark.parametrize("m,n", [(1, 1), (5, 3), (31, 33), (32, 32), (33, 31), (100, 70), (513, 129), (1000, 36), (2049, 64), (40, 1000)])
@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (0.75, -1.5)])
@pytest.mark.parametrize("pad", [0, 1, 4])
@pytest.mark.parametrize("cols", [1, 2])
def test_sgemv_t_unit_bit_exact(g, O, m, n, alpha, beta, pad, cols, monkeypatch):
    """pad 0/4 -> 16-byte cp.async path when n % 4 == 0; pad 1 -> 4-byte path; partial last row tile and column strip.
    cols = columns per lane (the library picks it from n; GUDA_GEMVT_COLS forces each instantiation; 2 uses packed FMUL2)."""
    monkeypatch.setenv("GUDA_GEMVT_COLS", str(cols))
    lda = n + pad
    a = O.uniform_f32(m * lda, 6); x = O.uniform_f32(m, 7); y = O.uniform_f32(n, 8)
    want = O.sgemv(True, m, n, alpha, a, lda, x, 1, beta, y.copy(), 1)
    got = gpu_gemv(g, True, m, n, alpha, a, lda, x, 1, beta, y, 1)
    assert np.array_equal(got, want), int(np.sum(got != want))


@pytest.mark.parametrize("trans", [False, True])
@pytest.mark.parametrize("incx,incy", [(2, 1), (1, 2), (-1, 1), (1, -1), (-2, -3), (3, -1), (7, 5)])
def test_sgemv_strided_bit_exact(g, O, trans, incx, incy):
    m, n = 67, 45
    lx, ly = (m, n) if trans else (n, m)
    a = O.uniform_f32(m * (n + 2), 9); x = O.uniform_f32(lx * abs(incx), 10); y = O.uniform_f32(ly * abs(incy), 11)
    for alpha, beta in [(1.0, 0.0), (-2.0, 0.5)]:
        want = O.sgemv(trans, m, n, alpha, a, n + 2, x, incx, beta, y.copy(), incy)
        got = gpu_gemv(g, trans, m, n, alpha, a, n + 2, x, incx, beta, y, incy)
        assert np.array_equal(got, want)


def test_sgemv_misaligned_base_pointers(g, O):
    """Sub-slices starting at odd element offsets (DevicePtr.Offset): GemvT must drop to its 4-byte staging path."""
    m, n, lda = 70, 64, 64
    a = O.uniform_f32(m * lda + 3, 12); x = O.uniform_f32(m + 1, 13); y = O.uniform_f32(n + 1, 14)
    for trans in (False, True):
        lx, ly = (m, n) if trans else (n, m)
        xs = O.uniform_f32(lx + 1, 15); ys = O.uniform_f32(ly + 1, 16)
        da, dx, dy = g.to_device(a), g.to_device(xs), g.to_device(ys)
        g.Sgemv(trans, m, n, 1.5, da.Offset(12), lda, dx.Offset(4), 1, 0.25, dy.Offset(4), 1)
        g.Synchronize()
        got = dy.Float32()
        want = ys.copy()
        O.sgemv(trans, m, n, 1.5, a[3:], lda, xs[1:], 1, 0.25, want[1:], 1)
        assert np.array_equal(got, want)
        g.Free(da); g.Free(dx); g.Free(dy)


def test_sgemv_special_values_and_quick_returns(g, O):
    m, n = 40, 70
    a = O.uniform_f32(m * n, 17); x = O.uniform_f32(n, 18)
    a[5 * n + 33] = np.nan; a[7 * n + 2] = np.inf; x[2] = 0.0
    y = O.uniform_f32(m, 19); y[3] = np.nan
    for alpha, beta in [(1.0, 0.0), (1.0, 2.0), (0.0, 0.0), (0.0, 1.0), (0.0, 3.0)]:
        want = O.sgemv(False, m, n, alpha, a, n, x, 1, beta, y.copy(), 1)
        got = gpu_gemv(g, False, m, n, alpha, a, n, x, 1, beta, y, 1)
        assert np.array_equal(got, want, equal_nan=True), (alpha, beta)
    xt = O.uniform_f32(m, 20); yt = O.uniform_f32(n, 21); yt[1] = np.inf
    for alpha, beta in [(1.0, 0.0), (1.0, 2.0), (0.0, 0.0), (0.0, 0.5)]:
        want = O.sgemv(True, m, n, alpha, a, n, xt, 1, beta, yt.copy(), 1)
        got = gpu_gemv(g, True, m, n, alpha, a, n, xt, 1, beta, yt, 1)
        assert np.array_equal(got, want, equal_nan=True), (alpha, beta)


def test_sgemv_t_beta_zero_clears_whole_y_slice(g, O):
    """gemv.go:61-64: `for i := range y { y[i] = 0 }` -- the host shim reproduces it (it knows len(y))."""
    m, n = 9, 5
    a = O.uniform_f32(m * n, 22); x = O.uniform_f32(m, 23); y = np.full(n + 6, 9.0, np.float32)
    want = O.sgemv(True, m, n, 1.0, a, n, x, 1, 0.0, y.copy(), 1)
    got = gpu_gemv(g, True, m, n, 1.0, a, n, x, 1, 0.0, y, 1)
    assert np.array_equal(got, want) and np.all(got[n:] == 0)


def test_sgemv_panics_match_reference_order(g):
    from guda_b200.api import BlasPanic
    a, x, y = g.to_device(np.ones(6, np.float32)), g.to_device(np.ones(3, np.float32)), g.to_device(np.ones(2, np.float32))
    cases = [((False, -1, 3, 1.0, a, 3, x, 1, 0.0, y, 1), "blas: m < 0"), ((False, 2, -1, 1.0, a, 3, x, 1, 0.0, y, 1), "blas: n < 0"),
             ((False, 2, 3, 1.0, a, 2, x, 1, 0.0, y, 1), "blas: bad leading dimension of A"),
             ((False, 2, 3, 1.0, a, 3, x, 0, 0.0, y, 1), "blas: zero x index increment"),
             ((False, 2, 3, 1.0, a, 3, x, 1, 0.0, y, 0), "blas: zero y index increment"),
             ((False, 2, 3, 1.0, a, 3, y, 1, 0.0, y, 1), "blas: insufficient length of x"),
             ((True, 2, 3, 1.0, a, 3, y, 1, 0.0, y, 1), "blas: insufficient length of y"),
             ((False, 2, 4, 1.0, a, 4, g.to_device(np.ones(4, np.float32)), 1, 0.0, y, 1), "blas: insufficient length of a")]
    for args, msg in cases:
        with pytest.raises(BlasPanic, match=msg):
            g.Sgemv(*args)
    with pytest.raises(BlasPanic, match="blas: insufficient length of y"):
        g.Sger(2, 3, 1.0, y, 1, y, 1, a, 3)
    with pytest.raises(BlasPanic, match="blas: bad leading dimension of A"):
        g.Sger(2, 3, 1.0, y, 1, x, 1, a, 2)
    g.Sgemv(False, 0, 3, 1.0, a, 3, x, 1, 0.0, y, 1)                  # m == 0: quick return before the length checks
    g.Synchronize()


@pytest.mark.parametrize("m,n,lda", [(1, 1, 1), (5, 3, 3), (33, 64, 64), (64, 128, 132), (70, 100, 101), (300, 36, 40), (31, 260, 260), (513, 1028, 1028)])
def test_sger_bit_exact(g, O, m, n, lda):
    a = O.uniform_f32(m * lda, 24); x = O.uniform_f32(m, 25); y = O.uniform_f32(n, 26)
    for alpha in (1.0, -0.3):
        want = O.sger(m, n, alpha, x, 1, y, 1, a.copy(), lda)
        got = gpu_ger(g, m, n, alpha, x, 1, y, 1, a, lda)
        assert np.array_equal(got, want)
    assert np.array_equal(gpu_ger(g, m, n, 0.0, x, 1, y, 1, a, lda), a)


def test_sgemv_large_properties(g, O):
    """4096 x 8192 (beyond what the row-by-row oracle is asked to do in full): sampled rows / columns against the oracle on the
    sub-problem, plus linearity in alpha by powers of two (exact in binary floating point)."""
    m, n = 4096, 8192
    a = O.uniform_f32(m * n, 27); x = O.uniform_f32(n, 28); xt = O.uniform_f32(m, 29)
    yN = gpu_gemv(g, False, m, n, 1.0, a, n, x, 1, 0.0, np.zeros(m, np.float32), 1)
    rows = [0, 1, 1000, 4095]
    sub = np.ascontiguousarray(a.reshape(m, n)[rows]).ravel()
    assert np.array_equal(yN[rows], O.sgemv(False, len(rows), n, 1.0, sub, n, x, 1, 0.0, np.zeros(len(rows), np.float32), 1))
    assert np.array_equal(gpu_gemv(g, False, m, n, 4.0, a, n, x, 1, 0.0, np.zeros(m, np.float32), 1), yN * np.float32(4))
    yT = gpu_gemv(g, True, m, n, 1.0, a, n, xt, 1, 0.0, np.zeros(n, np.float32), 1)
    cols = [0, 31, 32, 5000, 8191]
    subT = np.ascontiguousarray(a.reshape(m, n)[:, cols]).ravel()
    assert np.array_equal(yT[cols], O.sgemv(True, m, len(cols), 1.0, subT, len(cols), xt, 1, 0.0, np.zeros(len(cols), np.float32), 1))
    ref64 = a.reshape(m, n).astype(np.float64).T @ xt.astype(np.float64)
    assert np.allclose(yT, ref64, rtol=1e-3, atol=1e-2)


# ---------------------------------------------------------------- fp64 (Dgemv / Dger): bit-exact against the oracle
DGEMV = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "dgemv_cases.json")))["cases"]


def gpu_dgemv(g, trans, m, n, alpha, a, lda, x, incx, beta, y, incy):
    da, dx, dy = g.to_device(a), g.to_device(x), g.to_device(y)
    g.Dgemv(trans, m, n, alpha, da, lda, dx, incx, beta, dy, incy)
    g.Synchronize()
    out = dy.Float64()
    assert np.array_equal(da.Float64(), a) and np.array_equal(dx.Float64(), x)
    g.Free(da); g.Free(dx); g.Free(dy)
    return out


@pytest.mark.parametrize("case", DGEMV, ids=[c["Name"] for c in DGEMV])
def test_dgemv_testblas_cases(g, O, case):
    trans = "NoTrans" not in case["tA"]
    a = np.array(case["A"], np.float64).ravel()
    for sub in case["Subcases"]:
        incx = -case["incX"] if sub.get("mulXNeg1") else case["incX"]
        incy = -case["incY"] if sub.get("mulYNeg1") else case["incY"]
        x = np.array(case["x"], np.float64); y = np.array(case["y"], np.float64)
        got = gpu_dgemv(g, trans, case["m"], case["n"], sub["alpha"], a, case["n"], x, incx, sub["beta"], y, incy)
        assert np.array_equal(got, O.dgemv(trans, case["m"], case["n"], sub["alpha"], a, case["n"], x, incx, sub["beta"], y.copy(), incy))
        ans = np.array(sub["ans"], np.float64)
        assert np.all(np.abs(got - ans) <= 1e-14 * np.maximum(1.0, np.maximum(np.abs(got), np.abs(ans))))


@pytest.mark.parametrize("trans", [False, True])
@pytest.mark.parametrize("m,n", [(1, 1), (3, 5), (9, 4), (33, 67), (64, 64), (130, 1001), (1000, 37), (257, 4099),
                                 (8, 256), (264, 320), (519, 4098)])      # the last three: 16-byte aligned rows, the cp.async ring route of GemvN
@pytest.mark.parametrize("incx,incy", [(1, 1), (2, 1), (1, -3), (-1, 2)])
def test_dgemv_bit_exact(g, O, trans, m, n, incx, incy):
    lda = n + (m % 3)
    lx, ly = (m, n) if trans else (n, m)
    a = O.uniform_f64(m * lda, 41); x = O.uniform_f64((lx - 1) * abs(incx) + 1, 42); y = O.uniform_f64((ly - 1) * abs(incy) + 1, 43)
    for alpha, beta in [(1.0, 0.0), (0.75, -1.5), (2.0, 1.0)]:
        want = O.dgemv(trans, m, n, alpha, a, lda, x, incx, beta, y.copy(), incy)
        got = gpu_dgemv(g, trans, m, n, alpha, a, lda, x, incx, beta, y, incy)
        assert np.array_equal(got, want), (alpha, beta, int(np.sum(got != want)))


@pytest.mark.parametrize("m,n,lda", [(1, 1, 1), (5, 3, 3), (33, 64, 64), (70, 100, 101), (300, 36, 40), (513, 1028, 1028)])
@pytest.mark.parametrize("incx,incy", [(1, 1), (2, -1), (-3, 2)])
def test_dger_bit_exact(g, O, m, n, lda, incx, incy):
    a = O.uniform_f64(m * lda, 44); x = O.uniform_f64((m - 1) * abs(incx) + 1, 45); y = O.uniform_f64((n - 1) * abs(incy) + 1, 46)
    da, dx, dy = g.to_device(a), g.to_device(x), g.to_device(y)
    g.Dger(m, n, -0.3, dx, incx, dy, incy, da, lda)
    g.Synchronize()
    assert np.array_equal(da.Float64(), O.dger(m, n, -0.3, x, incx, y, incy, a.copy(), lda))
    g.Free(da); g.Free(dx); g.Free(dy)


def test_dgemv_quick_returns_and_panics(g, O):
    from guda_b200.api import BlasPanic
    a, x, y = g.to_device(np.ones(6)), g.to_device(np.ones(3)), g.to_device(np.full(2, 7.0))
    with pytest.raises(BlasPanic, match="blas: bad leading dimension of A"):
        g.Dgemv(False, 2, 3, 1.0, a, 2, x, 1, 0.0, y, 1)
    with pytest.raises(BlasPanic, match="blas: insufficient length of x"):
        g.Dgemv(False, 2, 3, 1.0, a, 3, y, 1, 0.0, y, 1)
    g.Dgemv(False, 2, 3, 0.0, a, 3, x, 1, 0.5, y, 1); g.Synchronize()
    assert np.array_equal(y.Float64(), np.full(2, 3.5))                    # alpha == 0: Dscal(beta)
    g.Dgemv(False, 2, 3, 1.0, a, 3, x, 1, 0.0, y, 1); g.Synchronize()
    assert np.array_equal(y.Float64(), np.full(2, 3.0))
