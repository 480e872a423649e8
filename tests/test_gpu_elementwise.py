"""GPU parity for the elementwise device ops and FusedKernel.Execute (guda_b200/csrc/elementwise.cu) through the C ABI:
bit for bit against oracle/elementwise.c (NaN compared as NaN)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def g():
    import guda_b200 as G
    G.Init(0)
    return G


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


def _specials(O, n, seed, lo=-12.0, hi=12.0):
    x = O.uniform_f32(n, seed, lo, hi)
    s = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 10.0, -10.0, 10.000001, -10.000001, 88.8, -87.4, 1e-45, -1e-45, 3.4e38], np.float32)
    k = min(n, s.size)
    x[:k] = s[:k]
    return x


def _same(a, b):
    """Same bits, except that any NaN equals any NaN (x86 and the GPU give 0 * Inf different sign / payload bits)."""
    nan = np.isnan(a)
    return np.array_equal(nan, np.isnan(b)) and np.array_equal(a[~nan].view(np.uint32), b[~nan].view(np.uint32))


@pytest.mark.parametrize("n", [1, 3, 4, 255, 1000, 4099, 1 << 20, (1 << 22) + 5])
def test_add_multiply_relu_sigmoid_bit_exact(g, O, n):
    a, b = _specials(O, n, 1 + n, -3e38, 3e38), _specials(O, n, 2 + n, -3e38, 3e38)[::-1].copy()
    da, db, dc = g.to_device(a), g.to_device(b), g.Malloc(n * 4)
    with np.errstate(all="ignore"):
        g.Add(da, db, dc, n); assert _same(dc.Float32()[:n], O.ew_add(a, b))
        g.Multiply(da, db, dc, n); assert _same(dc.Float32()[:n], O.ew_mul(a, b))
    x = _specials(O, n, 3 + n)
    dx = g.to_device(x); g.ReLU(dx, n); assert _same(dx.Float32()[:n], O.ew_relu(x)); g.Free(dx)
    dx = g.to_device(x); g.Sigmoid(dx, n); assert _same(dx.Float32()[:n], O.ew_sigmoid(x)); g.Free(dx)
    for p in (da, db, dc):
        g.Free(p)


def test_unaligned_views_and_untouched_tail(g, O):
    n = 1001
    x = _specials(O, n + 8, 5)
    dx = g.to_device(x)
    view = dx.Offset(4)                                   # 4-byte aligned only: the scalar route
    g.Sigmoid(view, n)
    got = dx.Float32()[:n + 8]
    assert got[0] == x[0] or (np.isnan(got[0]) and np.isnan(x[0]))
    assert _same(got[1:n + 1], O.ew_sigmoid(x[1:n + 1])) and _same(got[n + 1:], x[n + 1:])
    g.Free(dx)


def test_reference_fused_operations(g, O):
    """guda_test.go:193-240: ReLU(2*x + bias) through NewFusedKernel().MulScalar(2).Add(1).ReLU().Execute(x, [bias], N)."""
    N = 1000
    x, bias = O.uniform_f32(N, 11), O.uniform_f32(N, 12, 0.0, 0.1)
    dx, dbias = g.to_device(x), g.to_device(bias)
    g.NewFusedKernel().MulScalar(2.0).Add(1.0).ReLU().Execute(dx, [dbias], N)
    g.Synchronize()
    got = dx.Float32()[:N]
    assert np.all(np.abs(got - np.maximum(2.0 * x.astype(np.float64) + bias, 0)) <= 1e-5)          # the reference's check
    assert _same(got, O.fused_execute(x, [bias], [(3, 2.0, 0, 0), (0, 1.0, 0, 0), (4, 0.0, 0, 0)]))
    g.Free(dx); g.Free(dbias)


@pytest.mark.parametrize("n", [7, 4096, (1 << 20) + 3])
def test_fused_op_lists_bit_exact(g, O, n):
    x = _specials(O, n, 21, -6, 6)
    o1, o2 = O.uniform_f32(n, 22), O.uniform_f32(n, 23, 0.5, 1.5)
    d1, d2 = g.to_device(o1), g.to_device(o2)
    # every op type in one list; the third binary op has no operand left and is skipped
    dx = g.to_device(x)
    g.NewFusedKernel().Add(1.5).Multiply().AddScalar(-0.25).MulScalar(0.75).Sigmoid().Add(2.0).ReLU().Execute(dx, [d1, d2], n)
    want = O.fused_execute(x, [o1, o2], [(0, 1.5, 0, 0), (1, 0, 0, 0), (2, -0.25, 0, 0), (3, 0.75, 0, 0), (5, 0, 0, 0), (0, 2.0, 0, 0), (4, 0, 0, 0)])
    assert _same(dx.Float32()[:n], want)
    g.Free(dx)
    dx = g.to_device(x); g.LinearReLU(dx, 1.7, -0.3, n)
    assert _same(dx.Float32()[:n], O.fused_execute(x, [], [(3, 1.7, 0, 0), (2, -0.3, 0, 0), (4, 0, 0, 0)])); g.Free(dx)
    dx = g.to_device(x); g.GELU(dx, n)
    assert _same(dx.Float32()[:n], O.fused_execute(x, [], [(6, 0, 0, 0)])); g.Free(dx)
    g.Free(d1); g.Free(d2)


def test_fused_broadcast(g, O):
    rows, cols = 37, 24
    x = O.uniform_f32(rows * cols, 31)
    b = O.uniform_f32(cols, 32)
    dx, db = g.to_device(x), g.to_device(b)
    g.NewFusedKernel().AddBroadcast(0.5, 1).ReLU().Execute(dx, [db], rows * cols, [rows, cols])
    want = O.fused_execute(x, [b], [(0, 0.5, cols, cols), (4, 0, 0, 0)])          # fused.go:152-161: size = prod(shape[1:]), dim = shape[1]
    assert _same(dx.Float32()[:rows * cols], want)
    g.Free(dx); g.Free(db)


def test_errors(g):
    z = g.Malloc(64)
    with pytest.raises(g.ErrNotSupported):
        g.NewFusedKernel().Custom(lambda v: v)
    with pytest.raises(IndexError):
        g.Add(z, z, z, 17)
    fk = g.NewFusedKernel()
    for _ in range(9):
        fk.ReLU()
    with pytest.raises(g.GUDAError):
        fk.Execute(z, None, 16)
    g.ReLU(z, 0)
    g.Free(z)


@pytest.mark.parametrize("m,n", [(1, 1), (7, 5), (33, 129), (64, 128), (100, 1000), (1000, 100), (257, 515), (2048, 4096), (5, 100003)])
@pytest.mark.parametrize("axis", [0, 1])
def test_reduce_sum_max_axis_bit_exact(g, O, m, n, axis):
    x = O.uniform_f32(m * n, 41 + m + n, -100, 100)
    x[::37] = -0.0
    if m * n > 20:
        x[3] = np.nan; x[n if m > 1 else 5] = np.inf; x[-1] = -np.inf
    dx, dout = g.to_device(x), g.Malloc((n if axis == 0 else m) * 4)
    k = n if axis == 0 else m
    g.ReduceSum(dx, [m, n], axis, dout)
    with np.errstate(all="ignore"):
        assert _same(dout.Float32()[:k], O.reduce_axis(x, m, n, axis))
        g.ReduceMax(dx, [m, n], axis, dout)
        assert _same(dout.Float32()[:k], O.reduce_axis(x, m, n, axis, True))
    g.Free(dx); g.Free(dout)


def test_reduce_axis_unaligned_and_errors(g, O):
    m, n = 37, 130
    x = O.uniform_f32(m * n + 1, 77)
    dx, dout = g.to_device(x), g.Malloc(max(m, n) * 4)
    view = dx.Offset(4)
    g.ReduceSum(view, [m, n], 1, dout)
    assert _same(dout.Float32()[:m], O.reduce_axis(x[1:], m, n, 1))
    with pytest.raises(g.GUDAError) as e:
        g.ReduceSum(dx, [m, n], 2, dout)
    assert "invalid axis" in e.value.message
    with pytest.raises(g.GUDAError) as e:
        g.ReduceMax(dx, [m, n, 1], 0, dout)
    assert "invalid shape" in e.value.message
    g.Free(dx); g.Free(dout)


@pytest.mark.parametrize("n,k", [(1, 1), (7, 4), (100, 100), (1000, 10), (65537, 5), ((1 << 20) + 3, 16), (5000, 64)])
def test_topk_indices_bit_exact(g, O, n, k):
    x = O.uniform_f32(n, 91 + n)
    x = np.round(x * 8).astype(np.float32) / 8 if n >= 100 else x          # many exact ties: their order is the swap history's
    if n > 6:
        x[2] = np.nan; x[n // 2] = np.nan; x[5] = np.inf; x[n - 1] = np.inf
    dx = g.to_device(x)
    vals, idx = g.TopK(dx, n, k)
    wv, wi = O.topk(x, k)
    assert np.array_equal(idx, wi)
    assert _same(vals, wv)
    assert np.array_equal(dx.Float32()[:n], x, equal_nan=True)              # the input is not modified
    g.Free(dx)


def test_topk_leading_nan_and_errors(g, O):
    x = np.array([np.nan, 5, 1, 5, np.nan], np.float32)
    dx = g.to_device(x)
    vals, idx = g.TopK(dx.Offset(0), 5, 9)                                   # k clamped to n
    wv, wi = O.topk(x, 9)
    assert np.array_equal(idx, wi) and _same(vals, wv) and len(idx) == 5
    view = dx.Offset(4)                                                      # unaligned working copy start handled by the head path
    vals, idx = g.TopK(view, 4, 2)
    wv, wi = O.topk(x[1:], 2)
    assert np.array_equal(idx, wi) and _same(vals, wv)
    with pytest.raises(g.GUDAError) as e:
        g.TopK(dx, 5, 0)
    assert "invalid parameter" in e.value.message
    g.Free(dx)


def test_fast_division_sequences_agree_with_div_rn_on_every_float32_input(g):
    """geluFloat32 / SigmoidFloat32 are functions of one float: the kernels' shortened division sequences (gemm_simt.cuh div_const_rn,
    div_normal_rn) and the plain div.rn restatement are evaluated on all 2^32 bit patterns on the device; not one bit may differ."""
    import ctypes as C
    counts = (C.c_ulonglong * 2)()
    first = (C.c_uint * 2)()
    g.capi.check(g.load().guda_selftest_activations(counts, first), "SelftestActivations")
    assert (counts[0], counts[1]) == (0, 0), "GELU differs on %d inputs (first 0x%08x), Sigmoid on %d (first 0x%08x)" % (
        counts[0], first[0], counts[1], first[1])
