"""GPU parity: runtime API + Level-1 through the C ABI vs the CPU oracle (run with -m gpu on a B200)."""
import numpy as np
import pytest

from tests.helpers import TOL_REDUCE, d_slice_tol_equal, d_tol_equal, golden

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


# ------------------------------------------------------------------ runtime (guda_test.go:16-133,312-375)
def test_malloc_free_sizes(g):
    for size in (400, 4096, 1 << 20, 4 << 20):
        p = g.Malloc(size)
        assert p.ptr != 0 and p.Size() == size
        g.Free(p)


def test_memcpy_round_trip_h2d_d2d_d2h(g):
    n = 1024
    host = np.arange(n, dtype=np.float32) * np.float32(1.5)
    d1, d2 = g.Malloc(n * 4), g.Malloc(n * 4)
    g.Memcpy(d1, host, n * 4, g.MemcpyHostToDevice)
    g.Memcpy(d2, d1, n * 4, g.MemcpyDeviceToDevice)
    out = np.zeros(n, dtype=np.float32)
    g.Memcpy(out, d2, n * 4, g.MemcpyDeviceToHost)
    assert np.array_equal(out, host)
    assert np.array_equal(d2.Float32(), host)
    # typed views and Offset (memory.go:246-309)
    assert np.array_equal(d2.Offset(16).Float32(), host[4:])
    assert d2.Offset(16).Size() == n * 4 - 16
    g.Free(d1); g.Free(d2)


def test_double_free_and_unknown_pointer(g):
    p = g.Malloc(1024)
    g.Free(p)
    with pytest.raises(g.GUDAError, match="double free"):
        g.Free(p)
    with pytest.raises(g.GUDAError, match="not found"):
        g.Free(g.DevicePtr(p.ptr + 64, 10))          # Free(ptr.Offset(n)) fails in the reference too


def test_device_queries(g):
    assert g.GetDeviceCount() == 1                    # guda_test.go:332-335
    with pytest.raises(g.GUDAError):
        g.SetDevice(1)
    g.SetDevice(0)
    props = g.GetDeviceProperties(0)
    assert props.sm_count > 0 and props.cc_major >= 10
    with pytest.raises(g.GUDAError):
        g.GetDeviceProperties(3)


def test_pool_stats_bit_exact(g):
    # guda_test.go:339-375 + memory.go:150-227: 64-byte rounding, allocated falls on Free, peak never, first-fit reuse
    g.PoolTrim()                                       # empty free list, as in a fresh reference process
    a0, p0 = g.GetMemoryStats()
    x = g.Malloc(1000)                                 # -> 1024
    a1, p1 = g.GetMemoryStats()
    assert a1 - a0 == 1024 and p1 >= p0 and p1 >= a1
    y = g.Malloc(100)                                  # -> 128
    a2, p2 = g.GetMemoryStats()
    assert a2 - a1 == 128
    g.Free(x)
    a3, p3 = g.GetMemoryStats()
    assert a3 == a2 - 1024 and p3 == p2
    z = g.Malloc(500)                                  # first fit: reuses the 1024-byte block, counted at block size
    a4, _ = g.GetMemoryStats()
    assert z.ptr == x.ptr and a4 == a3 + 1024
    g.Free(y); g.Free(z)
    a5, p5 = g.GetMemoryStats()
    assert a5 == a0 and p5 >= p2


def test_streams_and_events(g):
    s = g.CreateStream()
    n = 1 << 20
    x = g.to_device(np.ones(n, dtype=np.float32)); y = g.to_device(np.zeros(n, dtype=np.float32))
    e0, e1 = g.Event(), g.Event()
    e0.Record(s)
    for _ in range(10):
        g.AXPY(1.0, x, y, n, stream=s)
    e1.Record(s)
    e1.Synchronize()
    assert e1.Query() and g.ElapsedTime(e0, e1) > 0
    s.Synchronize()
    assert s.Query()
    assert np.all(y.Float32() == 10.0)
    s2 = g.CreateStream()
    s2.WaitEvent(e1)
    g.Scale(0.5, y, n, stream=s2)
    s2.Synchronize()
    assert np.all(y.Float32() == 5.0)
    e0.Destroy(); e1.Destroy(); s.Destroy(); s2.Destroy()
    g.Free(x); g.Free(y)
    g.Synchronize()


def test_pool_reuse_is_ordered_after_work_queued_on_another_stream(g, O):
    """Free while a kernel on stream A still writes the block, Malloc again (first fit hands the same block back), use it on
    stream B: the pool must order the new owner after stream A's queued work (runtime.cu: an event recorded at Free, waited for
    at reuse).  The reference's synchronous BLAS never had this window (memory.go:207-227); an asynchronous drop-in does."""
    n = 4096
    sa, sb = g.CreateStream(), g.CreateStream()
    a, b = g.Malloc(n * n * 4), g.Malloc(n * n * 4)
    g.FillUniform(a, n, n, n, 1); g.FillUniform(b, n, n, n, 2)
    g.Synchronize()
    for _ in range(3):
        c = g.Malloc(n * n * 4 + 64)
        for _ in range(4):
            g.GEMM(False, False, n, n, n, 1.0, a, n, b, n, 0.0, c, n, stream=sa)      # ~9 ms of work writing c, queued on stream A
        g.Free(c)                                                                   # ... and freed under it
        q = g.Malloc(n * n * 4 + 64)
        assert q.ptr == c.ptr                                                       # first-fit reuse, as memory.go:160-180
        g.Memset(q, 0x5A, n * n * 4, stream=sb)                                     # the new owner runs on stream B
        sb.Synchronize(); sa.Synchronize()
        assert np.all(q.Byte()[:n * n * 4] == 0x5A)                                 # stream A's late writes did not land on top
        g.Free(q)
    for p in (a, b):
        g.Free(p)
    sa.Destroy(); sb.Destroy()
    g.PoolTrim()


def test_dot_device_results_on_many_streams_do_not_share_scratch(g, O):
    """12 streams x 6 device-result DOTs in flight at once: each stream owns its partials and ticket counter (level1.cu), so every
    result equals the deterministic single-stream value bit for bit."""
    n = (1 << 20) + 77
    L = g.load()
    xs = [g.to_device(O.uniform_f32(n, 300 + i)) for i in range(3)]
    ys = [g.to_device(O.uniform_f32(n, 400 + i)) for i in range(3)]
    want = [np.float32(g.DOT(xs[i], ys[i], n)) for i in range(3)]
    streams = [g.CreateStream() for _ in range(12)]
    outs = g.Malloc(12 * 6 * 4)
    for rep in range(6):
        for si, st in enumerate(streams):
            i = (si + rep) % 3
            g.capi.check(L.guda_sdot_device(n, xs[i].ptr, 1, ys[i].ptr, 1, outs.ptr + 4 * (si * 6 + rep), st.handle), "SdotDevice")
    for st in streams:
        st.Synchronize()
    got = outs.Float32()[:72].reshape(12, 6)
    for si in range(12):
        for rep in range(6):
            assert got[si, rep].view(np.uint32) == want[(si + rep) % 3].view(np.uint32), (si, rep)
    for st in streams:
        st.Destroy()
    for p in xs + ys + [outs]:
        g.Free(p)


# ------------------------------------------------------------------ Level 1 fp32
@pytest.mark.parametrize("n", [1, 3, 31, 32, 33, 1000, 1024, 4097, 1 << 20, (1 << 20) + 5])
def test_axpy_bit_exact(g, O, n):
    x = O.uniform_f32(n, 100 + n); y = O.uniform_f32(n, 200 + n)
    dx, dy = g.to_device(x), g.to_device(y)
    g.AXPY(2.5, dx, dy, n)                              # reference_test.go: alpha = 2.5
    g.Synchronize()
    want = O.saxpy(n, 2.5, x, 1, y.copy(), 1)
    assert np.array_equal(dy.Float32(), want)
    g.Free(dx); g.Free(dy)


def test_axpy_misaligned_views_bit_exact(g, O):
    n = 10000
    x = O.uniform_f32(n + 8, 1); y = O.uniform_f32(n + 8, 2)
    dx, dy = g.to_device(x), g.to_device(y)
    for ox, oy in [(1, 1), (1, 2), (3, 0), (0, 5)]:
        g.Memcpy(dy, y, y.nbytes, g.MemcpyHostToDevice)
        g.AXPY(-1.25, dx.Offset(4 * ox), dy.Offset(4 * oy), n)
        want = y.copy()
        O.saxpy(n, -1.25, x[ox:], 1, want[oy:], 1)
        assert np.array_equal(dy.Float32(), want), (ox, oy)
    g.Free(dx); g.Free(dy)


def test_axpy_alpha_zero_and_nan(g, O):
    n = 100
    x = np.full(n, np.nan, dtype=np.float32); y = O.uniform_f32(n, 3)
    dx, dy = g.to_device(x), g.to_device(y)
    g.AXPY(0.0, dx, dy, n)                              # alpha == 0 -> untouched (no NaN propagation)
    assert np.array_equal(dy.Float32(), y)
    g.AXPY(1.0, dx, dy, n)
    assert np.all(np.isnan(dy.Float32()))
    g.Free(dx); g.Free(dy)


@pytest.mark.parametrize("n", [1, 31, 32, 33, 1000, 1024, 65537, 1 << 22])
def test_dot_within_reduce_tolerance(g, O, n):
    x = O.uniform_f32(n, 7 + n); y = O.uniform_f32(n, 9 + n)
    dx, dy = g.to_device(x), g.to_device(y)
    got = g.DOT(dx, dy, n)
    want = O.sdot(n, x, 1, y, 1)
    exact = float(np.dot(x.astype(np.float64), y.astype(np.float64)))
    # ReduceArchTolerance, or -- when the reference's own order is the less accurate one and sits near that bound (n = 2^22: its
    # error against fp64 is 0.76 of the bound) -- no further from the oracle than twice the oracle's own distance to the fp64 value
    assert O.near_equal(float(want), float(got), TOL_REDUCE) or \
        abs(float(got) - float(want)) <= 2.0 * abs(float(want) - exact) + 4.0 * 2.0 ** -24 * abs(exact)
    assert abs(float(got) - exact) <= abs(float(want) - exact) + 1e-6 * max(1.0, abs(exact))   # at least as accurate as the reference order
    # deterministic
    assert g.DOT(dx, dy, n) == got
    g.Free(dx); g.Free(dy)


def test_baseline_kat_on_gpu(g):
    kat = golden("baseline_kat.json")["results"]
    i = np.arange(1000, dtype=np.float32)
    x = i * np.float32(0.001); y = i * np.float32(0.002)
    dx, dy = g.to_device(x), g.to_device(y)
    got = float(g.DOT(dx, dy, 1000))
    assert abs(got - kat["DOT_1000"]["checksum"]) <= 1e-4 * kat["DOT_1000"]["checksum"]
    g.Free(dx); g.Free(dy)


@pytest.mark.parametrize("n", [1, 5, 1000, 4099, 1 << 20])
def test_scale_bit_exact(g, O, n):
    x = O.uniform_f32(n, 17 + n)
    dx = g.to_device(x)
    g.Scale(1.7, dx, n)
    assert np.array_equal(dx.Float32(), O.sscal(n, 1.7, x.copy(), 1))
    xn = x.copy(); xn[::3] = np.nan; xn[1::7] = np.inf
    g.Memcpy(dx, xn, xn.nbytes, g.MemcpyHostToDevice)
    g.Scale(0.0, dx, n)                                  # alpha == 0 -> exact zero fill, NaN/Inf cleared
    assert np.array_equal(dx.Float32(), np.zeros(n, dtype=np.float32))
    g.Free(dx)


def test_level1_fp32_increments_and_panics(g, O):
    for c in golden("level1_double.json")["two_vector"]:
        x = np.array(c["X"], dtype=np.float32); y = np.array(c["Y"], dtype=np.float32)
        dx, dy = g.to_device(x), g.to_device(y)
        dx.size, dy.size = x.nbytes, y.nbytes
        if c["Panic"]:
            with pytest.raises(g.BlasPanic):
                g.Sdot(c["N"], dx, c["Incx"], dy, c["Incy"])
            with pytest.raises(g.BlasPanic):
                g.Saxpy(c["N"], 2.0, dx, c["Incx"], dy, c["Incy"])
        else:
            assert float(g.Sdot(c["N"], dx, c["Incx"], dy, c["Incy"])) == float(c.get("DdotAns", 0)), c["Name"]
            for kind in c["DaxpyCases"]:
                g.Memcpy(dy, y, y.nbytes, g.MemcpyHostToDevice)
                g.Saxpy(c["N"], kind["Alpha"], dx, c["Incx"], dy, c["Incy"])
                assert np.array_equal(dy.Float32(), np.array(kind["Ans"], dtype=np.float32)), c["Name"]
        g.Free(dx); g.Free(dy)


def test_saxpy_strided_unfused_bit_exact(g, O):
    n = 5000
    x = O.uniform_f32(3 * n, 41); y = O.uniform_f32(2 * n, 42)
    dx, dy = g.to_device(x), g.to_device(y)
    for incx, incy in [(3, 2), (-3, 2), (2, -1), (-1, -2)]:
        g.Memcpy(dy, y, y.nbytes, g.MemcpyHostToDevice)
        g.Saxpy(n, 0.3, dx, incx, dy, incy)
        want = O.saxpy(n, 0.3, x, incx, y.copy(), incy)
        assert np.array_equal(dy.Float32(), want), (incx, incy)
    g.Free(dx); g.Free(dy)


# ------------------------------------------------------------------ Level 1 fp64 (testblas tables + oracle)
def test_testblas_daxpy_ddot_on_gpu(g):
    for c in golden("level1_double.json")["two_vector"]:
        x = np.array(c["X"], dtype=np.float64); y = np.array(c["Y"], dtype=np.float64)
        dx, dy = g.to_device(x), g.to_device(y)
        dx.size, dy.size = x.nbytes, y.nbytes
        if c["Panic"]:
            with pytest.raises(g.BlasPanic):
                g.Ddot(c["N"], dx, c["Incx"], dy, c["Incy"])
            with pytest.raises(g.BlasPanic):
                g.Daxpy(c["N"], 1.0, dx, c["Incx"], dy, c["Incy"])
        else:
            assert d_tol_equal(g.Ddot(c["N"], dx, c["Incx"], dy, c["Incy"]), float(c.get("DdotAns", 0))), c["Name"]
            for kind in c["DaxpyCases"]:
                g.Memcpy(dy, y, y.nbytes, g.MemcpyHostToDevice)
                g.Daxpy(c["N"], kind["Alpha"], dx, c["Incx"], dy, c["Incy"])
                assert d_slice_tol_equal(dy.Float64(), kind["Ans"]), c["Name"]
        g.Free(dx); g.Free(dy)


def test_testblas_dscal_on_gpu(g):
    for c in golden("level1_double.json")["one_vector"]:
        for kind in c.get("DscalCases", []):
            x = np.array(c["X"], dtype=np.float64)
            dx = g.to_device(x)
            dx.size = x.nbytes
            if c["Panic"]:
                with pytest.raises(g.BlasPanic):
                    g.Dscal(c["N"], kind["Alpha"], dx, c["Incx"])
            else:
                g.Dscal(c["N"], kind["Alpha"], dx, c["Incx"])
                assert d_slice_tol_equal(dx.Float64(), kind["Ans"]), c["Name"]
            g.Free(dx)


@pytest.mark.parametrize("n,off", [(1000, 0), (1003, 0), (1003, 1), (4098, 3), (17, 2)])
def test_daxpy_alignment_rule_bit_exact(g, O, n, off):
    # f64 AxpyUnitary: unfused peel to 32-byte alignment of y, FMA body, unfused last (n-peel)%4
    x = O.uniform_f64(n + 4, 5); y = O.uniform_f64(n + 4, 6)
    dx, dy = g.to_device(x), g.to_device(y)
    g.Daxpy(n, 0.7, dx.Offset(8 * off), 1, dy.Offset(8 * off), 1)
    base = np.zeros(n + 8, dtype=np.float64)
    start = (-base.ctypes.data // 8) % 4                 # make the host copy 32-byte aligned like cudaMalloc memory
    yh = base[start:start + n + 4]
    yh[:] = y
    O.daxpy(n, 0.7, x[off:], 1, yh[off:], 1)
    assert np.array_equal(dy.Float64(), yh)
    g.Free(dx); g.Free(dy)


def test_ddot_large(g, O):
    n = 1 << 21
    x = O.uniform_f64(n, 1); y = O.uniform_f64(n, 2)
    dx, dy = g.to_device(x), g.to_device(y)
    got = g.Ddot(n, dx, 1, dy, 1)
    want = O.ddot(n, x, 1, y, 1)
    assert abs(got - want) <= 1e-12 * max(1.0, abs(want))
    g.Free(dx); g.Free(dy)
