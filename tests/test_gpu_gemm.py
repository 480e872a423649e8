"""GPU parity: SGEMM / DGEMM / fused / bf16 / fp16 through the C ABI vs the CPU oracle (run with -m gpu on a B200)."""
import numpy as np
import pytest

from tests.helpers import TOL_GEMM, d_slice_tol_equal, flatten, gemm_parity, golden, transpose

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


def _run_sgemm(g, ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc):
    da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c)
    g.GEMM(bool(ta), bool(tb), m, n, k, alpha, da, lda, db, ldb, beta, dc, ldc)
    g.Synchronize()
    out = dc.Float32()
    assert np.array_equal(da.Float32(), a, equal_nan=True) and np.array_equal(db.Float32(), b, equal_nan=True)     # inputs untouched
    g.Free(da); g.Free(db); g.Free(dc)
    return out


SHAPES = [(1, 1, 1), (8, 8, 8), (16, 16, 16), (32, 32, 32), (63, 65, 64), (128, 128, 128), (100, 200, 50),
          (257, 129, 77), (64, 64, 1), (5, 300, 33), (256, 256, 256), (1024, 1024, 64)]


@pytest.mark.parametrize("m,n,k", SHAPES)
@pytest.mark.parametrize("ta,tb", [(0, 0), (1, 0), (1, 1)])
def test_sgemm_bit_exact_nn_tn_tt(g, O, m, n, k, ta, tb):
    """NN/TN follow the AxpyUnitary FMA chain, TT the unfused AxpyInc chain: bit-for-bit equal to the oracle."""
    lda = (m if ta else k) + 3
    ldb = (k if tb else n) + 1
    ldc = n + 2
    a = O.uniform_f32((k if ta else m) * lda, 1); b = O.uniform_f32((n if tb else k) * ldb, 2)
    c0 = O.uniform_f32(m * ldc, 3, 0.0, 1.0)
    got = _run_sgemm(g, ta, tb, m, n, k, 1.5, a, lda, b, ldb, 0.5, c0, ldc)
    want = O.sgemm(ta, tb, m, n, k, 1.5, a, lda, b, ldb, 0.5, c0.copy(), ldc, threads=8)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("alpha", [np.inf, -np.inf])
@pytest.mark.parametrize("m,n,k", [(65, 67, 1), (130, 40, 2), (7, 5, 3), (65, 67, 19)])
def test_sgemm_infinite_alpha_on_ragged_shapes(g, O, m, n, k, alpha):
    """alpha = +-Inf on shapes that take the edge path (zero padding beyond m / n / k): the padding must stay an exact zero --
    0 * Inf = NaN would pass the `tmp != 0` test and poison whole tiles; the reference never visits those indices."""
    a = O.uniform_f32(m * k, 5); b = O.uniform_f32(k * n, 6); c0 = O.uniform_f32(m * n, 7)
    got = _run_sgemm(g, 0, 0, m, n, k, float(alpha), a, k, b, n, 0.5, c0, n)
    want = O.sgemm(0, 0, m, n, k, float(alpha), a, k, b, n, 0.5, c0.copy(), n)
    assert np.array_equal(got, want, equal_nan=True)
    assert not np.isfinite(want).any()
    if k == 1:
        assert not np.isnan(want).any()          # one term per element: +-Inf everywhere, a NaN could only come from the padding


def test_gemm_bf16_exact_route_keeps_negative_zero_on_ragged_k(g, O):
    """bfloat16.go:136-157 has no zero test: sum = beta*c; sum += (alpha*a)*b for the REAL k only.  With a = 0, b < 0 every product
    is -0 and a -0 start stays -0; one extra padded step (+0 * +0) would flip it to +0."""
    m, n, k = 5, 9, 3
    a = O.to_bf16_fast(np.zeros(m * k, np.float32))
    b = O.to_bf16_fast(-np.abs(O.uniform_f32(k * n, 8)) - np.float32(0.5))
    c0 = O.to_bf16_fast(np.full(m * n, -0.0, np.float32))
    da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
    g.SetGemm16Mode(1)
    try:
        g.GEMMBFloat16(False, False, m, n, k, 1.0, da, k, db, n, 1.0, dc, n)
        g.Synchronize()
    finally:
        g.SetGemm16Mode(0)
    want = O.gemm_bf16(m, n, k, 1.0, a, k, b, n, 1.0, c0.copy(), n)
    assert np.array_equal(dc.UInt16(), want) and np.all(want == 0x8000)
    for p in (da, db, dc):
        g.Free(p)


@pytest.mark.parametrize("m,n,k", [(9, 7, 8), (63, 65, 64), (128, 128, 128), (257, 129, 77), (512, 512, 300)])
def test_sgemm_nt_within_tolerance(g, O, m, n, k):
    """NotTrans x Trans is the DotUnitary form (32-lane blocked order): compared under the K-scaled GEMM tolerance.
    (Contiguous 4^3/8^3/16^3 with bTrans is excluded on purpose: compute/sgemm.go:98-111 multiplies by B as stored --
    a reference defect SURVEY.md a7 says not to replicate; test_sgemm_small_btrans_is_true_transpose pins our answer.)"""
    a = O.uniform_f32(m * k, 1); b = O.uniform_f32(n * k, 2); c0 = O.uniform_f32(m * n, 3, 0.0, 1.0)
    got = _run_sgemm(g, 0, 1, m, n, k, 1.5, a, k, b, k, 0.5, c0, n)
    want = O.sgemm(0, 1, m, n, k, 1.5, a, k, b, k, 0.5, c0.copy(), n, threads=8)
    ab = 1.5 * np.abs(a.reshape(m, k).astype(np.float64)) @ np.abs(b.reshape(n, k).astype(np.float64)).T
    nfail, v = gemm_parity(want, got, TOL_GEMM, ab, k, c_abs=0.5 * np.abs(c0))
    assert nfail == 0, v


@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (1.0, 1.0), (0.0, 0.5), (2.0, 0.0), (0.0, 1.0)])
def test_sgemm_alpha_beta_forms(g, O, alpha, beta):
    m, n, k = 130, 70, 90
    a = O.uniform_f32(m * k, 4); b = O.uniform_f32(k * n, 5); c0 = O.uniform_f32(m * n, 6)
    got = _run_sgemm(g, 0, 0, m, n, k, alpha, a, k, b, n, beta, c0, n)
    want = O.sgemm(0, 0, m, n, k, alpha, a, k, b, n, beta, c0.copy(), n)
    assert np.array_equal(got, want)


def test_sgemm_beta_zero_clears_nan_and_zero_skip(g, O):
    m, n, k = 96, 80, 48
    a = O.uniform_f32(m * k, 7); b = O.uniform_f32(k * n, 8)
    a[::5] = 0.0                                    # zero entries of A are skipped (compute/sgemm.go:301) ...
    b[3 * n + 7] = np.inf                           # ... so an Inf in B only poisons rows whose a[i,3] != 0
    b[10 * n + 1] = np.nan
    c0 = np.full(m * n, np.nan, dtype=np.float32)   # beta == 0 -> C zeroed first, NaNs do not propagate
    got = _run_sgemm(g, 0, 0, m, n, k, 1.0, a, k, b, n, 0.0, c0, n)
    want = O.sgemm(0, 0, m, n, k, 1.0, a, k, b, n, 0.0, c0.copy(), n)
    assert np.array_equal(got, want, equal_nan=True)
    assert np.isnan(want).sum() < m * n and np.isnan(want).sum() > 0


@pytest.mark.parametrize("s", [4, 8, 16])
def test_sgemm_small_fast_paths_bit_exact(g, O, s):
    a = O.uniform_f32(s * s, 21); b = O.uniform_f32(s * s, 22); c0 = O.uniform_f32(s * s, 23)
    for beta in (0.5, 0.0):
        got = _run_sgemm(g, 0, 0, s, s, s, 1.5, a, s, b, s, beta, c0, s)
        want = O.sgemm(0, 0, s, s, s, 1.5, a, s, b, s, beta, c0.copy(), s)
        assert np.array_equal(got, want)


def test_sgemm_small_btrans_is_true_transpose(g, O):
    s = 8
    a = O.uniform_f32(s * s, 21); b = O.uniform_f32(s * s, 22); c0 = np.zeros(s * s, np.float32)
    got = _run_sgemm(g, 0, 1, s, s, s, 1.0, a, s, b, s, 0.0, c0, s)
    want = a.reshape(s, s).astype(np.float64) @ b.reshape(s, s).astype(np.float64).T
    assert np.max(np.abs(got.reshape(s, s) - want)) <= 1e-5


def test_sgemm_panics(g):
    d = g.Malloc(64)
    with pytest.raises(g.BlasPanic, match="m < 0"):
        g.GEMM(False, False, -1, 4, 4, 1.0, d, 4, d, 4, 0.0, d, 4)
    with pytest.raises(g.BlasPanic, match="bad leading dimension of A"):
        g.GEMM(False, False, 4, 4, 4, 1.0, d, 3, d, 4, 0.0, d, 4)
    with pytest.raises(g.BlasPanic, match="bad leading dimension of B"):
        g.GEMM(False, True, 4, 4, 4, 1.0, d, 4, d, 3, 0.0, d, 4)
    with pytest.raises(g.BlasPanic, match="bad leading dimension of C"):
        g.GEMM(False, False, 4, 4, 4, 1.0, d, 4, d, 4, 0.0, d, 3)
    with pytest.raises(g.BlasPanic, match="insufficient length of a"):
        g.GEMM(False, False, 5, 4, 4, 1.0, d, 4, d, 4, 0.0, d, 4)
    g.GEMM(False, False, 0, 4, 4, 1.0, d, 4, d, 4, 0.0, d, 4)       # m == 0 -> quick return
    g.Free(d)


def test_hilbert_and_reference_property(g, O):
    n = 100                                           # math_correctness_test.go:328-381
    h = np.array([[1.0 / (i + j + 1) for j in range(n)] for i in range(n)], dtype=np.float32)
    got = _run_sgemm(g, 0, 0, n, n, n, 1.0, h.ravel(), n, np.eye(n, dtype=np.float32).ravel(), n, 0.0, np.zeros(n * n, np.float32), n)
    assert np.array_equal(got.reshape(n, n), h)
    # reference_test.go: GEMM 32^3 alpha=1.5 beta=0.5 vs scalar Reference, 1e-5 abs-or-rel
    a = O.uniform_f32(32 * 32, 1); b = O.uniform_f32(32 * 32, 2); c0 = O.uniform_f32(32 * 32, 3)
    got = _run_sgemm(g, 0, 0, 32, 32, 32, 1.5, a, 32, b, 32, 0.5, c0, 32)
    ref = O.ref_gemm(0, 0, 32, 32, 32, 1.5, a, 32, b, 32, 0.5, c0.copy(), 32)
    d = np.abs(got - ref)
    assert np.all((d <= 1e-5) | (d <= 1e-5 * np.abs(ref)))


def test_sgemm_1024_full_compare(g, O):
    """BASELINE config 1 (README headline size): full elementwise bit compare with the oracle."""
    n = 1024
    a = O.uniform_f32(n * n, 11); b = O.uniform_f32(n * n, 12); c0 = O.uniform_f32(n * n, 13, 0.0, 1.0)
    got = _run_sgemm(g, 0, 0, n, n, n, 1.5, a, n, b, n, 0.5, c0, n)
    want = O.sgemm(0, 0, n, n, n, 1.5, a, n, b, n, 0.5, c0.copy(), n, threads=8)
    assert np.array_equal(got, want)


# ------------------------------------------------------------------ DGEMM
@pytest.mark.parametrize("case_idx", range(7))
@pytest.mark.parametrize("ta,tb", [(0, 0), (1, 0), (0, 1), (1, 1)])
def test_testblas_dgemm_on_gpu(g, case_idx, ta, tb):
    case = golden("dgemm_cases.json")["cases"][case_idx]
    a_rows = transpose(case["a"]) if ta else case["a"]
    b_rows = transpose(case["b"]) if tb else case["b"]
    a, b, c = flatten(a_rows), flatten(b_rows), flatten(case["c"])
    da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c)
    g.Dgemm(bool(ta), bool(tb), case["m"], case["n"], case["k"], case["alpha"], da, len(a_rows[0]), db, len(b_rows[0]),
            case["beta"], dc, len(case["c"][0]))
    g.Synchronize()
    assert np.array_equal(da.Float64(), a) and np.array_equal(db.Float64(), b)
    assert d_slice_tol_equal(dc.Float64(), flatten(case["ans"]))
    g.Free(da); g.Free(db); g.Free(dc)


@pytest.mark.parametrize("m,n,k", [(64, 64, 64), (128, 128, 96), (67, 35, 29), (130, 66, 40), (256, 512, 128), (3, 2, 5)])
@pytest.mark.parametrize("ta,tb,alpha", [(0, 0, 1.0), (0, 0, 2.0), (1, 0, 2.0), (1, 1, 2.0)])
def test_dgemm_bit_exact(g, O, m, n, k, ta, tb, alpha):
    """alpha==1 NN: 8x8/4x4 FMA tiles + unfused edges; alpha!=1 NN and TN: fp64 AxpyUnitary (alignment rule); TT: unfused."""
    lda = m if ta else k
    ldb = k if tb else n
    a = O.uniform_f64((k if ta else m) * lda, 1); b = O.uniform_f64((n if tb else k) * ldb, 2); c0 = O.uniform_f64(m * n, 3)
    da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
    g.Dgemm(bool(ta), bool(tb), m, n, k, alpha, da, lda, db, ldb, 0.5, dc, n)
    g.Synchronize()
    base = np.zeros(m * n + 8, dtype=np.float64)
    start = (-base.ctypes.data // 8) % 4              # host C 32-byte aligned, like device memory
    ch = base[start:start + m * n]
    ch[:] = c0
    O.dgemm(ta, tb, m, n, k, alpha, a, lda, b, ldb, 0.5, ch, n)
    assert np.array_equal(dc.Float64(), ch)
    g.Free(da); g.Free(db); g.Free(dc)


def test_dgemm_nt_tolerance(g, O):
    m, n, k = 130, 70, 200
    a = O.uniform_f64(m * k, 1); b = O.uniform_f64(n * k, 2); c0 = O.uniform_f64(m * n, 3)
    da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
    g.Dgemm(False, True, m, n, k, 2.0, da, k, db, k, 0.5, dc, n)
    want = O.dgemm(0, 1, m, n, k, 2.0, a, k, b, k, 0.5, c0.copy(), n)
    assert np.max(np.abs(dc.Float64() - want)) <= 1e-13
    g.Free(da); g.Free(db); g.Free(dc)


# ------------------------------------------------------------------ fused GEMM + bias + activation
def _fused_inputs():
    m, n, k = 8, 8, 4
    a = np.full((m, k), 0.1, dtype=np.float32)
    for i in range(min(m, k)):
        a[i, i] = 1.0
    b = np.tile(np.arange(n, dtype=np.float32) * np.float32(0.1), (k, 1))
    bias = np.arange(n, dtype=np.float32) * np.float32(0.05)
    return m, n, k, a.ravel().copy(), b.ravel().copy(), bias


def test_fused_gelu_kat(g, O):
    # fused_gelu_test.go:81-170 (tolerance 0.015 there; the scalar route is reproduced bit-for-bit here)
    m, n, k, a, b, bias = _fused_inputs()
    da, db, dbias, dc = g.to_device(a), g.to_device(b), g.to_device(bias), g.Malloc(m * n * 4)
    g.FusedGEMMBiasGELU(False, False, m, n, k, 1.0, da, k, db, n, 0.0, dc, n, dbias)
    g.Synchronize()
    want = O.fused_gemm_bias_gelu(0, 0, m, n, k, 1.0, a, k, b, n, 0.0, np.zeros(m * n, np.float32), n, bias)
    assert np.array_equal(dc.Float32(), want)
    pade = O.fused_gemm_bias_gelu(0, 0, m, n, k, 1.0, a, k, b, n, 0.0, np.zeros(m * n, np.float32), n, bias, route="pade")
    assert np.max(np.abs(dc.Float32() - pade)) <= 0.015
    with pytest.raises(g.ErrNotSupported):
        g.FusedGEMMBiasGELU(True, False, m, n, k, 1.0, da, k, db, n, 0.0, dc, n, dbias)
    with pytest.raises(g.ErrNotSupported):
        g.FusedGEMMBiasReLU(False, False, m, n, k, 1.0, da, k, db, n, 0.5, dc, n, dbias)
    for p in (da, db, dbias, dc):
        g.Free(p)


@pytest.mark.parametrize("sz,alpha", [(8, 1.0), (16, 2.0), (32, 1.0), (64, 1.5), (200, 1.0), (512, 0.7)])
@pytest.mark.parametrize("act", ["relu", "gelu"])
def test_fused_bias_act_bit_exact(g, O, sz, alpha, act):
    # fused_ops_test.go:15-90 sizes/alphas (+ two larger ones); GEMM part = stock FMA chain, epilogue in registers
    m, n, k = sz, sz + (8 if sz >= 200 else 0), sz
    a = O.uniform_f32(m * k, 1 + sz); b = O.uniform_f32(k * n, 2 + sz); bias = O.uniform_f32(n, 3 + sz)
    da, db, dbias, dc = g.to_device(a), g.to_device(b), g.to_device(bias), g.Malloc(m * n * 4)
    if act == "relu":
        g.FusedGEMMBiasReLU(False, False, m, n, k, alpha, da, k, db, n, 0.0, dc, n, dbias)
        want = O.fused_gemm_bias_relu(0, 0, m, n, k, alpha, a, k, b, n, 0.0, np.zeros(m * n, np.float32), n, bias, threads=8)
    else:
        g.FusedGEMMBiasGELU(False, False, m, n, k, alpha, da, k, db, n, 0.0, dc, n, dbias)
        want = O.fused_gemm_bias_gelu(0, 0, m, n, k, alpha, a, k, b, n, 0.0, np.zeros(m * n, np.float32), n, bias, threads=8)
    g.Synchronize()
    assert np.array_equal(dc.Float32(), want)
    for p in (da, db, dbias, dc):
        g.Free(p)


def test_add_bias_gelu_kat(g, O):
    x = ((np.arange(16) - 8).astype(np.float32) * np.float32(0.25))      # fused_gelu_test.go:36-78
    bias = np.arange(4).astype(np.float32) * np.float32(0.1)
    dx, dbias = g.to_device(x), g.to_device(bias)
    g.AddBiasGELU(dx, dbias, 16)
    assert np.array_equal(dx.Float32(), O.add_bias_gelu(x.copy(), bias))
    g.Free(dx); g.Free(dbias)


def test_gelu_exact_over_a_sweep(g, O):
    x = np.concatenate([np.linspace(-12, 12, 20001, dtype=np.float32), O.uniform_f32(20000, 5, -6, 6),
                        np.array([0.0, -0.0, 0.5, -0.5, 10.0, -10.0, 10.5, 1e-20, 88.0, -88.0, np.inf, -np.inf], np.float32)])
    dx, dz = g.to_device(x), g.to_device(np.zeros(1, np.float32))
    g.AddBiasGELU(dx, dz, x.size)
    want = O.gelu(x)
    assert np.array_equal(dx.Float32(), want, equal_nan=True)
    g.Free(dx); g.Free(dz)


# ------------------------------------------------------------------ bf16 / fp16
def test_conversions_match_reference_rules(g, O):
    x = np.concatenate([O.uniform_f32(5000, 3, -70000, 70000), O.uniform_f32(5000, 4, -1e-4, 1e-4),
                        np.array([0x3F817FFF, 0x3F807FFF, 0x3F81FFFF], dtype=np.uint32).view(np.float32)])
    import ctypes as C
    L = g.load()
    dx = g.to_device(x); d16 = g.Malloc(x.size * 2); d32 = g.Malloc(x.size * 4)
    assert L.guda_f32_to_bf16(dx.ptr, d16.ptr, x.size, None) == 0
    assert np.array_equal(d16.UInt16(), O.to_bf16(x))
    assert L.guda_bf16_to_f32(d16.ptr, d32.ptr, x.size, None) == 0
    assert np.array_equal(d32.Float32(), O.bf16_to_f32(O.to_bf16(x)))
    assert L.guda_f32_to_f16(dx.ptr, d16.ptr, x.size, None) == 0
    assert np.array_equal(d16.UInt16(), O.f32_to_f16(x))
    h = np.arange(0, 65536, 3, dtype=np.uint16)
    dh = g.to_device(h); df = g.Malloc(h.size * 4)
    assert L.guda_f16_to_f32(dh.ptr, df.ptr, h.size, None) == 0
    assert np.array_equal(df.Float32().view(np.uint32), O.f16_to_f32(h).view(np.uint32))
    for p in (dx, d16, d32, dh, df):
        g.Free(p)


@pytest.mark.parametrize("m,n,k", [(8, 8, 8), (64, 64, 64), (70, 50, 33), (128, 256, 96)])
def test_gemm_bf16_exact_route_bit_exact(g, O, m, n, k):
    a = O.to_bf16_fast(O.uniform_f32(m * k, 1)); b = O.to_bf16_fast(O.uniform_f32(k * n, 2)); c0 = O.to_bf16_fast(O.uniform_f32(m * n, 3))
    g.SetGemm16Mode(1)
    try:
        for alpha, beta in [(1.0, 0.0), (1.5, 0.5)]:
            da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
            g.GEMMBFloat16(False, False, m, n, k, alpha, da, k, db, n, beta, dc, n)
            want = O.gemm_bf16(m, n, k, alpha, a, k, b, n, beta, c0.copy(), n)
            assert np.array_equal(dc.UInt16(), want)
            g.Free(da); g.Free(db); g.Free(dc)
    finally:
        g.SetGemm16Mode(0)


@pytest.mark.parametrize("m,n,k", [(8, 8, 8), (16, 16, 16), (64, 48, 40), (130, 70, 90)])
def test_gemm_f16_exact_route_bit_exact(g, O, m, n, k):
    a = O.f32_to_f16(O.uniform_f32(m * k, 1)); b = O.f32_to_f16(O.uniform_f32(k * n, 2)); c0 = O.f32_to_f16(O.uniform_f32(m * n, 3))
    g.SetGemm16Mode(1)
    try:
        da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
        g.GEMMFloat16(False, False, m, n, k, 1.5, da, k, db, n, 0.5, dc, n)
        want = O.gemm_f16(0, 0, m, n, k, 1.5, a, k, b, n, 0.5, c0.copy(), n, threads=4)
        assert np.array_equal(dc.UInt16(), want)
        g.Free(da); g.Free(db); g.Free(dc)
    finally:
        g.SetGemm16Mode(0)


# ------------------------------------------------------------------ K-blocked launches, fused all-gather building block
def test_sgemm_k_blocked_launches_bit_exact(g, O):
    """K >= 16384 with a large C runs as 8192-deep launches that resume the chain from the stored fp32 accumulator
    (gemm.cu launch_simt): the bits must equal the oracle's single chain."""
    m, n, k = 2048, 2048, 16500
    a = O.uniform_f32(m * k, 71); b = O.uniform_f32(k * n, 72); c0 = O.uniform_f32(m * n, 73)
    got = _run_sgemm(g, 0, 0, m, n, k, 1.5, a, k, b, n, 0.5, c0, n)
    want = O.sgemm(0, 0, m, n, k, 1.5, a, k, b, n, 0.5, c0.copy(), n, threads=16)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("pr,pc,m,n,k", [(2, 2, 256, 384, 100), (1, 2, 128, 520, 64), (2, 4, 512, 1024, 33)])
def test_sgemm_scatter_places_blocks_like_an_all_gather(g, O, pr, pc, m, n, k):
    """guda_sgemm_scatter is the kernel behind the fused GEMM + all-gather: every emulated rank computes its block and its
    epilogue stores it into EVERY rank's gathered matrix.  With the ranks emulated on one GPU (one call per rank, P
    destination buffers) each buffer must end up equal to the single-GPU product, bit for bit."""
    import ctypes as C
    from guda_b200 import sharding
    L = g.load()
    world = pr * pc
    a = O.uniform_f32(m * k, 81); b = O.uniform_f32(k * n, 82)
    da, db = g.to_device(a), g.to_device(b)
    fulls = [g.Malloc(m * n * 4) for _ in range(world)]
    for f in fulls:
        g.Memset(f, 0xFF, m * n * 4)
    dests = (C.c_void_p * world)(*[f.ptr for f in fulls])
    blocks = []
    for rank in range(world):
        r0, rows, c0, cols = sharding.block_extents(rank, pr, pc, m, n)
        blk = g.Malloc(rows * cols * 4)
        blocks.append((blk, r0, rows, c0, cols))
        g.capi.check(L.guda_sgemm_scatter(0, 0, rows, cols, k, 1.25, da.ptr + r0 * k * 4, k, db.ptr + c0 * 4, n, 0.0, blk.ptr, cols,
                                          dests, world, n, r0, c0, None), "SgemmScatter")
    g.Synchronize()
    want = O.sgemm(0, 0, m, n, k, 1.25, a, k, b, n, 0.0, np.zeros(m * n, np.float32), n)
    for f in fulls:
        assert np.array_equal(f.Float32(), want)
    for blk, r0, rows, c0, cols in blocks:
        assert np.array_equal(blk.Float32().reshape(rows, cols), want.reshape(m, n)[r0:r0 + rows, c0:c0 + cols])
    # c may be NULL when only the gathered copies are wanted
    g.Memset(fulls[0], 0, m * n * 4)
    one = (C.c_void_p * 1)(fulls[0].ptr)
    g.capi.check(L.guda_sgemm_scatter(0, 0, m, n, k, 1.25, da.ptr, k, db.ptr, n, 0.0, None, n, one, 1, n, 0, 0, None), "SgemmScatter")
    g.Synchronize()
    assert np.array_equal(fulls[0].Float32(), want)
    for p in [da, db] + fulls + [b_[0] for b_ in blocks]:
        g.Free(p)
