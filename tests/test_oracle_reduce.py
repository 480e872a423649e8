"""The reductions oracle (oracle/reduce.c) against the reference's own tables (reduce_test.go:28-240, tolerance 1e-5 as
floatEquals) and against the semantics of the AVX2 kernels it restates (VMAXPS operand rule, VCOMISS + JBE)."""
import json
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reduce_cases.json")))["cases"]
FN = {"testSum": O.ssum, "testMax": O.smax, "testMin": O.smin, "testArgMax": O.sargmax, "testArgMin": O.sargmin, "testSumSquares": O.ssumsq}


def _exp(e):
    return {"-inf": -np.inf, "inf": np.inf}.get(e, e)


@pytest.mark.parametrize("fn", sorted(GOLD))
def test_reference_tables(fn):
    for row in GOLD[fn]:
        x = np.asarray(row["input"], np.float32)
        got = FN[fn](x)
        want = _exp(row["expected"])
        if np.isinf(want):
            assert got == want, (fn, row["name"])
        else:
            assert abs(float(got) - want) <= 1e-5, (fn, row["name"], got, want)


def _fold(x, is_max):
    """Pure-Python replay of MaxAVX2 / MinAVX2 (reduce_avx2_amd64.s:9-113)."""
    op = (lambda a, b: a if a > b else b) if is_max else (lambda a, b: a if a < b else b)
    n = len(x)
    if n == 0:
        return np.float32(-np.inf if is_max else np.inf)
    r = x[0]
    i = 0
    if n >= 8:
        acc = [x[0]] * 8
        for g in range(n // 8):
            for l in range(8):
                acc[l] = op(acc[l], x[8 * g + l])
        q = [op(acc[l], acc[l + 4]) for l in range(4)]
        q = [op(q[0], q[2]), op(q[1], q[3])]
        r = op(q[0], q[1])
        i = 8 * (n // 8)
    for j in range(i if n >= 8 else 0, n):
        r = op(r, x[j])
    return r


@pytest.mark.parametrize("n", [1, 3, 7, 8, 9, 15, 16, 17, 64, 100, 1001])
def test_minmax_matches_lane_replay_with_nans_and_signed_zeros(n):
    rng = np.random.default_rng(n)
    for trial in range(20):
        x = rng.standard_normal(n).astype(np.float32)
        if trial % 2:
            x[rng.integers(0, n, max(1, n // 5))] = np.nan
        if trial % 3 == 0:
            x[:] = np.where(rng.random(n) < 0.5, np.float32(0.0), np.float32(-0.0))
        for is_max, f in ((True, O.smax), (False, O.smin)):
            got, want = f(x), _fold(list(x), is_max)
            assert (np.isnan(got) and np.isnan(want)) or (got == want and np.signbit(got) == np.signbit(want)), (n, trial, is_max)


def test_arg_semantics():
    nan = np.nan
    assert O.sargmax(np.array([nan, 5, 6], np.float32)) == 0            # NaN x[0] never loses (VCOMISS unordered -> JBE)
    assert O.sargmax(np.array([1, nan, 6, nan, 6], np.float32)) == 2    # NaNs skipped, first of equal maxima
    assert O.sargmin(np.array([1, nan, -6, nan, -6], np.float32)) == 2
    assert O.sargmax(np.array([0.0, -0.0, 0.0], np.float32)) == 0       # +-0 compare equal
    assert O.sargmax(np.zeros(0, np.float32)) == -1 and O.sargmin(np.zeros(0, np.float32)) == -1
    assert O.isamax(5, np.array([1, -7, 7, nan, 3], np.float32), 1) == 1
    assert O.isamax(3, np.array([1, 9, -7, 9, 8], np.float32), 2) == 2  # strided: elements 1, -7, 8 -> index 2
    assert O.isamax(0, np.zeros(0, np.float32), 1) == -1 and O.isamax(3, np.ones(3, np.float32), -1) == -1
    with pytest.raises(O.OraclePanic, match="blas: zero x index increment"):
        O.isamax(3, np.ones(3, np.float32), 0)
    with pytest.raises(O.OraclePanic, match="blas: insufficient length of x"):
        O.isamax(3, np.ones(4, np.float32), 2)
    with pytest.raises(O.OraclePanic, match="blas: n < 0"):
        O.isamax(-1, np.ones(4, np.float32), 1)


def test_sums_follow_the_lane_orders():
    x = O.uniform_f32(1000, 5)
    assert abs(float(O.ssum(x)) - float(np.sum(x.astype(np.float64)))) < 1e-3
    assert abs(float(O.ssumsq(x)) - float(np.sum(x.astype(np.float64) ** 2))) < 1e-3
    # SumSquares, n = 8: 8 FMA lanes, (l + l+4), hadd, hadd
    y = np.arange(1, 9, dtype=np.float32) * np.float32(1.1)
    sq = [np.float32(np.float64(v) * np.float64(v)) for v in y]           # fma(x, x, 0) = fl(x*x)
    q = [np.float32(sq[l] + sq[l + 4]) for l in range(4)]
    want = np.float32(np.float32(q[0] + q[1]) + np.float32(q[2] + q[3]))
    assert O.ssumsq(y) == want


def test_softmax_reference_properties_and_fast_exp():
    """reduce_test.go:297-333 (sum == 1 within 1e-5, values in [0, 1]) and fastExp against exp within its design error."""
    for inp in ([1, 1, 1, 1], [1, 2, 3, 4], [10, 20, 30, 40]):
        x = O.softmax(np.asarray(inp, np.float32))
        assert abs(float(O.ssum(x)) - 1.0) <= 1e-5 and np.all((x >= 0) & (x <= 1))
        want = np.exp(np.asarray(inp, np.float64) - max(inp)); want /= want.sum()
        assert np.allclose(x, want, rtol=2e-5, atol=1e-7)
    for v in (0.0, -0.1, -1.0, -5.5, -20.0, -87.0, -100.0):
        assert abs(float(O.fast_exp(v)) - np.exp(v)) <= 2e-6 * np.exp(v) + 1e-38


def test_layer_norm_matches_float64_model():
    x = O.uniform_f32(256, 21); gamma = O.uniform_f32(256, 22); beta = O.uniform_f32(256, 23)
    got = O.layer_norm(x.copy(), gamma, beta, 1e-5)
    x64 = x.astype(np.float64)
    want = (x64 - x64.mean()) / np.sqrt(x64.var() + 1e-5) * gamma + beta
    assert np.allclose(got, want, rtol=1e-4, atol=1e-4)                 # reference_test.go:338: tol * 10
    plain = O.layer_norm(x.copy(), None, beta, 1e-5)                    # gamma or beta missing: no affine step (fused.go:299)
    assert np.allclose(plain, (x64 - x64.mean()) / np.sqrt(x64.var() + 1e-5), rtol=1e-4, atol=1e-4)


def test_dsdot_sdsdot_oracle():
    """compute/level1float32_dsdot.go / _sdsdot.go: exact products, double accumulation; panics in the reference's order."""
    for n, incx, incy in [(1, 1, 1), (7, 1, 1), (8, 1, 1), (1001, 1, 1), (300, 3, -2), (100, -1, 2)]:
        x = O.uniform_f32(n * abs(incx) + 1, 5); y = O.uniform_f32(n * abs(incy) + 1, 6)
        xs = x[:(n - 1) * abs(incx) + 1:abs(incx)][:n].astype(np.float64); ys = y[:(n - 1) * abs(incy) + 1:abs(incy)][:n].astype(np.float64)
        if incx < 0: xs = xs[::-1]
        if incy < 0: ys = ys[::-1]
        want = float(np.dot(xs, ys))
        got = O.dsdot(n, x, incx, y, incy)
        assert abs(got - want) <= 1e-13 * max(1.0, float(np.sum(np.abs(xs * ys))))
        assert O.sdsdot(n, 2.5, x, incx, y, incy) == np.float32(np.float32(2.5) + np.float32(got))
    one = np.ones(4, np.float32)
    assert O.dsdot(0, one, 1, one, 1) == 0.0 and O.sdsdot(0, 2.5, one, 1, one, 1) == 0.0
    for args, msg in [((3, one, 0, one, 1), "blas: zero x index increment"), ((3, one, 1, one, 0), "blas: zero y index increment"),
                      ((-1, one, 1, one, 1), "blas: n < 0"), ((5, one, 1, one, 1), "blas: insufficient length of x"),
                      ((3, one, 1, one, 2), "blas: insufficient length of y")]:
        with pytest.raises(O.OraclePanic, match=msg):
            O.dsdot(*args)
