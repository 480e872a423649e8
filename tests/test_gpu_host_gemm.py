"""guda_sgemm_host (guda_b200/csrc/host_gemm.cu): SGEMM on host-resident operands with the copies pipelined under the kernel.
Must be bit-identical to guda_sgemm on device-resident copies (K-chunked accumulation resumes the same FMA chain) and to the oracle."""
import ctypes as C

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


def device_gemm(g, ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c0, ldc):
    da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
    g.Sgemm(bool(ta), bool(tb), m, n, k, alpha, da, lda, db, ldb, beta, dc, ldc)
    g.Synchronize()
    out = dc.Float32()
    g.Free(da); g.Free(db); g.Free(dc)
    return out


@pytest.mark.parametrize("ta,tb", [(0, 0), (1, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("m,n,k", [(1, 1, 1), (16, 16, 16), (67, 45, 33), (300, 260, 257), (2304, 128, 64)])
@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (0.5, -1.5)])
def test_host_gemm_matches_oracle_and_device_path(g, O, ta, tb, m, n, k, alpha, beta):
    lda = (m if ta else k) + 3
    ldb = (k if tb else n) + 1
    ldc = n + 2
    a = O.uniform_f32((k if ta else m) * lda, 1); b = O.uniform_f32((n if tb else k) * ldb, 2); c0 = O.uniform_f32(m * ldc, 3)
    c = c0.copy()
    g.SgemmHost(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc)
    dev = device_gemm(g, ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c0, ldc)
    assert np.array_equal(c.reshape(m, ldc)[:, :n], dev.reshape(m, ldc)[:, :n])            # same bits as the device-resident call
    if not (tb and not ta):                      # NT is the DotUnitary form: tolerance-matched on the device path too
        want = O.sgemm(ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c0.copy(), ldc)
        assert np.array_equal(c.reshape(m, ldc)[:, :n], want.reshape(m, ldc)[:, :n])
    assert np.array_equal(c.reshape(m, ldc)[:, n:], c0.reshape(m, ldc)[:, n:])          # padding columns of C untouched


@pytest.mark.parametrize("n", [4, 8, 16])
def test_host_gemm_small_cube_paths_follow_contiguity(g, O, n):
    """compute/sgemm.go:80-96: the 4^3/8^3/16^3 kernels (own summation order) apply to contiguous operands only."""
    a = O.uniform_f32(n * n, 7); b = O.uniform_f32(n * n, 8); c0 = O.uniform_f32(n * n, 9)
    c = c0.copy()
    g.SgemmHost(0, 0, n, n, n, 1.0, a, n, b, n, 0.0, c, n)
    assert np.array_equal(c, O.sgemm(0, 0, n, n, n, 1.0, a, n, b, n, 0.0, c0.copy(), n))
    ap = O.uniform_f32(n * (n + 1), 7); cp = O.uniform_f32(n * (n + 1), 9)
    c = cp.copy()
    g.SgemmHost(0, 0, n, n, n, 1.0, ap, n + 1, b, n, 0.0, c, n + 1)
    assert np.array_equal(c, O.sgemm(0, 0, n, n, n, 1.0, ap, n + 1, b, n, 0.0, cp.copy(), n + 1))


def test_host_gemm_k_chunked_pipeline_same_bits(g, O):
    """k = 3 chunks of 8192 (+ a ragged one) and 8 row panels: the pipelined path itself.  Pinned host memory, as bench.py uses it."""
    m, n, k = 2304, 640, 3 * 8192 + 100
    a = O.uniform_f32(m * k, 4); b = O.uniform_f32(k * n, 5); c0 = O.uniform_f32(m * n, 6)
    ref = device_gemm(g, 0, 0, m, n, k, 1.0, a, k, b, n, 0.5, c0, n)
    L = g.load()
    ha, hb, hc = C.c_void_p(), C.c_void_p(), C.c_void_p()
    for h, arr in ((ha, a), (hb, b), (hc, c0)):
        g.capi.check(L.guda_malloc_host(C.byref(h), arr.nbytes), "MallocHost")
        C.memmove(h, arr.ctypes.data, arr.nbytes)
    g.SgemmHost(0, 0, m, n, k, 1.0, ha.value, k, hb.value, n, 0.5, hc.value, n)
    got = np.ctypeslib.as_array(C.cast(hc, C.POINTER(C.c_float)), shape=(m * n,)).copy()
    for h in (ha, hb, hc):
        L.guda_free_host(h)
    assert np.array_equal(got, ref)
    rows = [0, 1, 1151, 2303]                                   # oracle on a row subset (the full 2304 x 640 x 24676 takes too long)
    sub_a = np.ascontiguousarray(a.reshape(m, k)[rows]).ravel(); sub_c = np.ascontiguousarray(c0.reshape(m, n)[rows]).ravel()
    want = O.sgemm(0, 0, len(rows), n, k, 1.0, sub_a, k, b, n, 0.5, sub_c, n)
    assert np.array_equal(got.reshape(m, n)[rows].ravel(), want)


def test_host_gemm_panics_and_quick_returns(g):
    from guda_b200.api import BlasPanic
    a = np.ones(6, np.float32); b = np.ones(6, np.float32); c = np.full(4, 7.0, np.float32)
    with pytest.raises(BlasPanic, match="blas: m < 0"):
        g.SgemmHost(0, 0, -1, 2, 3, 1.0, a, 3, b, 2, 0.0, c, 2)
    with pytest.raises(BlasPanic, match="blas: bad leading dimension of B"):
        g.SgemmHost(0, 0, 2, 2, 3, 1.0, a, 3, b, 1, 0.0, c, 2)
    with pytest.raises(BlasPanic, match="blas: insufficient length of a"):
        g.SgemmHost(0, 0, 2, 2, 3, 1.0, a[:5], 3, b, 2, 0.0, c, 2)
    g.SgemmHost(0, 0, 0, 2, 3, 1.0, a, 3, b, 2, 0.0, c, 2)
    assert np.all(c == 7.0)
    g.SgemmHost(0, 0, 2, 2, 0, 1.0, a, 1, b, 2, 0.5, c, 2)       # k == 0: C = beta * C
    assert np.all(c == 3.5)
    g.SgemmHost(0, 0, 2, 2, 3, 1.0, a, 3, b, 2, 0.0, c, 2)
    assert np.all(c == 3.0)
