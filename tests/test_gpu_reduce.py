"""GPU parity for the reductions (guda_b200/csrc/reduce.cu) through the C ABI.  Index outputs, Max and Min are bit-exact against
the oracle's replay of reduce_avx2_amd64.s (NaN and signed-zero behaviour included); Sum / SumSquares are checked under
ReduceArchTolerance (tolerance_arch.go:99-106: abs 1e-5, rel 1e-4, 16 ULP) scaled like the reference's own DOT checks."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reduce_cases.json")))["cases"]


@pytest.fixture(scope="module")
def g():
    import guda_b200 as G
    G.Init(0)
    return G


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


def same_bits(a, b):
    return (np.isnan(a) and np.isnan(b)) or (a == b and np.signbit(a) == np.signbit(b))


def test_reference_tables(g, O):
    meth = {"testSum": "Sum", "testMax": "Max", "testMin": "Min", "testArgMax": "ArgMax", "testArgMin": "ArgMin", "testSumSquares": "SumSquares"}
    for fn, rows in GOLD.items():
        for row in rows:
            x = np.asarray(row["input"], np.float32)
            d = g.to_device(x) if x.size else g.Malloc(4)
            if not x.size:
                d.size = 0
            got = getattr(d, meth[fn])(x.size)
            want = {"-inf": -np.inf, "inf": np.inf}.get(row["expected"], row["expected"])
            if isinstance(want, float) and np.isinf(want):
                assert got == want
            else:
                assert abs(float(got) - want) <= 1e-5, (fn, row["name"], got, want)
            g.Free(d)


SIZES = [1, 2, 7, 8, 9, 31, 100, 1000, 4099, 65536, 1 << 20, (1 << 20) + 3, 5_000_001]


@pytest.mark.parametrize("n", SIZES)
def test_arg_and_extremes_bit_exact(g, O, n):
    x = O.uniform_f32(n, 100 + n % 97)
    d = g.to_device(x)
    assert d.ArgMax(n) == O.sargmax(x) and d.ArgMin(n) == O.sargmin(x)
    assert g.Isamax(n, d, 1) == O.isamax(n, x, 1)
    assert same_bits(d.Max(n), O.smax(x)) and same_bits(d.Min(n), O.smin(x))
    g.Free(d)


@pytest.mark.parametrize("n", [5, 8, 64, 1000, 4099, 1 << 18, 3_000_001])
def test_ties_pick_the_first_index(g, O, n):
    rng = np.random.default_rng(n)
    x = rng.integers(-3, 4, n).astype(np.float32)          # many equal extremes
    d = g.to_device(x)
    assert d.ArgMax(n) == O.sargmax(x) == int(np.argmax(x))
    assert d.ArgMin(n) == O.sargmin(x) == int(np.argmin(x))
    assert g.Isamax(n, d, 1) == O.isamax(n, x, 1) == int(np.argmax(np.abs(x)))
    g.Free(d)


@pytest.mark.parametrize("n", [1, 7, 8, 9, 17, 100, 1001, 70001, 1 << 20])
def test_nan_and_signed_zero_follow_the_avx2_kernels(g, O, n):
    rng = np.random.default_rng(7 * n + 1)
    for trial in range(8):
        x = rng.standard_normal(n).astype(np.float32)
        if trial in (0, 1, 2):
            x[rng.integers(0, n, max(1, n // 7))] = np.nan
        if trial == 1:
            x[0] = np.nan
        if trial == 2:
            x[-1] = np.nan
        if trial == 3:
            x[:] = np.where(rng.random(n) < 0.5, np.float32(0.0), np.float32(-0.0))        # extreme is a signed zero
        if trial == 4:
            x[:] = np.nan
        if trial == 5:
            x = -np.abs(x); x[rng.integers(0, n, max(1, n // 3))] = np.where(rng.random(max(1, n // 3)) < 0.5, 0.0, -0.0)   # max is +-0
        if trial == 6:
            x = np.maximum(x, np.float32(0.0))               # post-ReLU: the Min is +0 and no -0 exists -> no replay needed, same bits
        if trial == 7:
            x = -np.abs(x); x[rng.integers(0, n, max(1, n // 3))] = np.float32(-0.0)          # the Max is -0 and no +0 exists
        d = g.to_device(x)
        assert same_bits(d.Max(n), O.smax(x)), (n, trial)
        assert same_bits(d.Min(n), O.smin(x)), (n, trial)
        assert d.ArgMax(n) == O.sargmax(x) and d.ArgMin(n) == O.sargmin(x), (n, trial)
        assert g.Isamax(n, d, 1) == O.isamax(n, x, 1), (n, trial)
        g.Free(d)


@pytest.mark.parametrize("n", SIZES)
def test_sum_and_sumsquares_within_reduce_tolerance(g, O, n):
    x = O.uniform_f32(n, 200 + n % 89)
    d = g.to_device(x)
    exact, exact2 = float(np.sum(x.astype(np.float64))), float(np.sum(x.astype(np.float64) ** 2))
    scale = float(np.sum(np.abs(x.astype(np.float64))))
    for got, ref, want, sc in ((d.Sum(n), O.ssum(x), exact, scale), (d.SumSquares(n), O.ssumsq(x), exact2, exact2)):
        # ReduceArchTolerance relative bound against the reference's own result, with the float64 sum as the arbiter: the GPU's
        # pairwise tree must be at least as close to it as 1e-4 relative (it is far closer than the 16/8-lane reference order)
        assert abs(float(got) - want) <= 1e-5 + 1e-4 * abs(want) + 1e-6 * sc, (n, got, want)
        # ... and within the same bound of the reference's result, allowing for the reference's own distance from the true sum
        assert abs(float(got) - float(ref)) <= 1e-5 + 1e-4 * abs(float(ref)) + 2e-6 * sc + abs(float(ref) - want)
    assert d.Sum(n) == d.Sum(n)                              # deterministic
    assert abs(float(d.Mean(n)) - exact / n) <= 1e-5 + 1e-4 * abs(exact / n) + 1e-6 * sc / n
    g.Free(d)


def test_offsets_strides_and_edges(g, O):
    from guda_b200.api import BlasPanic
    x = O.uniform_f32(10007, 3)
    d = g.to_device(x)
    for off in (1, 2, 3, 5):                                 # misaligned starts (DevicePtr.Offset)
        v = x[off:]
        s = d.Offset(4 * off)
        assert s.ArgMax(v.size) == O.sargmax(v) and s.ArgMin(v.size) == O.sargmin(v)
        assert same_bits(s.Max(v.size), O.smax(v)) and same_bits(s.Min(v.size), O.smin(v))
        assert abs(float(s.Sum(v.size)) - float(np.sum(v.astype(np.float64)))) < 1e-2
    for inc in (2, 3, 7):
        n = (x.size - 1) // inc + 1
        assert g.Isamax(n, d, inc) == O.isamax(n, x, inc)
    assert g.Isamax(0, d, 1) == -1 and g.Isamax(1, d, 1) == 0 and g.Isamax(5, d, -1) == -1
    with pytest.raises(BlasPanic, match="blas: zero x index increment"):
        g.Isamax(5, d, 0)
    with pytest.raises(BlasPanic, match="blas: insufficient length of x"):
        g.Isamax(x.size, d, 2)
    with pytest.raises(BlasPanic, match="blas: n < 0"):
        g.Isamax(-1, d, 1)
    assert d.ArgMax(0) == -1 and d.ArgMin(0) == -1 and d.Max(0) == -np.inf and d.Min(0) == np.inf and d.Sum(0) == 0 and d.SumSquares(0) == 0
    with pytest.raises(IndexError):
        d.Sum(x.size + 1)
    g.Free(d)


def test_large_2_pow_26(g, O):
    """BASELINE-scale vector (2^26 elements, 256 MiB): indices against numpy on the host copy, extremes bit-exact."""
    n = 1 << 26
    d = g.Malloc(n * 4)
    g.FillUniform(d, 1, n, n, 77)
    x = d.Float32()
    assert d.ArgMax(n) == int(np.argmax(x)) and d.ArgMin(n) == int(np.argmin(x))
    assert g.Isamax(n, d, 1) == int(np.argmax(np.abs(x)))
    assert d.Max(n) == x.max() and d.Min(n) == x.min()
    assert abs(float(d.Sum(n)) - float(np.sum(x.astype(np.float64)))) <= 1e-4 * float(np.sum(np.abs(x.astype(np.float64)))) * 1e-2
    g.Free(d)


@pytest.mark.parametrize("n", [1, 4, 5, 1000, 4099, 1 << 20, 3_000_001])
def test_softmax_against_oracle(g, O, n):
    """exp step bit-exact given the (bit-exact) max; the normaliser differs only by summation order -> a few ULP."""
    x = (O.uniform_f32(n, 300 + n % 7) * np.float32(8.0)).astype(np.float32)
    want = O.softmax(x.copy())
    d = g.to_device(x)
    g.Softmax(d, n)
    got = d.Float32()
    assert np.all((got >= 0) & (got <= 1)) and abs(float(np.sum(got.astype(np.float64))) - 1.0) <= 1e-5     # reduce_test.go:318-329
    # got_i = e_i * fl(1/S_gpu), want_i = e_i * fl(1/S_ref) with identical e_i: the ratio is ONE constant for every i (to the two
    # roundings of the scale step), and it is near 1 up to the reference's own sequential-sum error (~n/2 ulp)
    nz = want > 1e-30
    ratio = got[nz].astype(np.float64) / want[nz].astype(np.float64)
    assert ratio.max() - ratio.min() <= 3e-7 * ratio.mean()
    assert abs(ratio.mean() - 1.0) <= 4e-6 + 1e-9 * n
    assert int(np.argmax(got)) == int(np.argmax(want))
    g.Free(d)


def test_softmax_reference_cases(g):
    for inp in ([1, 1, 1, 1], [1, 2, 3, 4], [10, 20, 30, 40]):
        d = g.to_device(np.asarray(inp, np.float32))
        g.Softmax(d, 4)
        assert abs(float(d.Sum(4)) - 1.0) <= 1e-5
        g.Free(d)


@pytest.mark.parametrize("n", [1, 2, 256, 1000, 4099, 1 << 20])
@pytest.mark.parametrize("affine", [True, False])
def test_layer_norm_against_oracle(g, O, n, affine):
    x = O.uniform_f32(n, 400 + n % 5); gamma = O.uniform_f32(n, 401); beta = O.uniform_f32(n, 402)
    want = O.layer_norm(x.copy(), gamma if affine else None, beta if affine else None, 1e-5)
    dx, dg, db = g.to_device(x), g.to_device(gamma), g.to_device(beta)
    g.LayerNorm(dx, dg if affine else None, db if affine else None, n, 1e-5)
    got = dx.Float32()
    x64 = x.astype(np.float64)
    model = (x64 - x64.mean()) / np.sqrt(x64.var() + 1e-5)
    if affine:
        model = model * gamma + beta
    assert np.allclose(got, model, rtol=1e-4, atol=1e-4)                # reference_test.go:338 (tol * 10), float64 model as arbiter
    ref_err = float(np.max(np.abs(want - model)))                       # the reference's own sequential-sum error (grows with n)
    assert float(np.max(np.abs(got - want))) <= 1e-4 + 2.0 * ref_err
    assert np.array_equal(dg.Float32(), gamma) and np.array_equal(db.Float32(), beta)
    g.Free(dx); g.Free(dg); g.Free(db)


@pytest.mark.parametrize("n,incx,incy", [(1, 1, 1), (7, 1, 1), (1001, 1, 1), (1 << 20, 1, 1), (3_000_001, 1, 1), (300, 3, -2), (5000, -1, 2), (4099, 2, 2)])
def test_dsdot_sdsdot(g, O, n, incx, incy):
    """fp32 inputs, fp64 accumulation: products are exact, so GPU and reference differ only by the order of the double adds."""
    from guda_b200.api import BlasPanic
    x = O.uniform_f32((n - 1) * abs(incx) + 1, 500 + n % 11); y = O.uniform_f32((n - 1) * abs(incy) + 1, 501)
    dx, dy = g.to_device(x), g.to_device(y)
    want = O.dsdot(n, x, incx, y, incy)
    got = g.Dsdot(n, dx, incx, dy, incy)
    # the reference adds ~n/4 doubles per lane sequentially (error up to ~n/4 ulp of the running sum); the GPU tree is tighter.
    # Bound both against a float64 pairwise sum of the exact products.
    xs, ys = x[::abs(incx)][:n].astype(np.float64), y[::abs(incy)][:n].astype(np.float64)
    if (incx < 0) != (incy < 0):
        ys = ys[::-1]
    exact, mag = float(np.sum(xs * ys)), float(np.sum(np.abs(xs * ys)))
    assert abs(got - exact) <= 1e-14 * mag + 1e-15
    assert abs(got - want) <= 1e-14 * mag + abs(want - exact) + 1e-15
    s = g.Sdsdot(n, 2.5, dx, incx, dy, incy)
    assert abs(float(s) - float(O.sdsdot(n, 2.5, x, incx, y, incy))) <= 2.4e-7 * max(1.0, abs(float(s)))       # one fp32 ulp
    assert g.Dsdot(0, dx, incx, dy, incy) == 0.0 and g.Sdsdot(0, 2.5, dx, incx, dy, incy) == 0.0
    with pytest.raises(BlasPanic, match="blas: n < 0"):
        g.Dsdot(-1, dx, incx, dy, incy)
    with pytest.raises(BlasPanic, match="blas: insufficient length of x"):
        g.Dsdot(n + 1, dx, incx, dy, incy)
    g.Free(dx); g.Free(dy)
