"""GPU parity for the tcgen05 tensor-core routes (bf16 / fp16 storage GEMM, fused epilogue, 3xTF32 SGEMM).

These routes change only the fp32 summation ORDER (and, for 3xTF32, split each fp32 product in three TF32 products),
so they are compared with the oracle under stated tolerances rather than bit-for-bit:
  bf16 / fp16 : |got - want| <= 1 ULP of the 16-bit format (2^-7 resp. 2^-10 relative) -- the reference truncates, so a
                sum that lands within ~1e-6 of a truncation boundary may fall on either side
  3xTF32      : |got - want| <= 8 * K * 2^-24 * sum_l |a_il||b_lj|  (the K-scaled bound of SURVEY 8(d) with the measured
                split constant), and the error against an fp64 product is reported
"""
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


def _bf16_vals(O, n, seed, lo=-1.0, hi=1.0):
    return O.to_bf16_fast(O.uniform_f32(n, seed, lo, hi))


def _tol(op):
    """The sm_100a entry of the tolerance framework (guda.GetOperationTolerance(op)): HalfULPTol / KScale are the stated bounds."""
    from guda_b200 import tolerance as T
    return T.GetOperationTolerance(op)


def _ulp16_check(got_f, want_f, mant_bits, order_bound=0.0, op="gemm_bf16"):
    """HalfULPTol ULPs of the 16-bit format, plus the fp32 summation-order bound (matters only for sums that cancel to ~0)."""
    from guda_b200 import tolerance as T
    ulp = T.half_ulp_bound(_tol(op), want_f, mant_bits, -14)
    bad = np.abs(got_f.astype(np.float64) - want_f.astype(np.float64)) > ulp * 1.001 + order_bound
    return int(bad.sum()), float(np.mean(got_f == want_f))


def _order_bound(O, a16, lda, b16, ldb, m, n, k, alpha, beta, c16, conv, op="gemm_bf16"):
    from guda_b200 import tolerance as T
    A = np.abs(conv(a16).reshape(-1)[:m * lda].reshape(m, lda)[:, :k].astype(np.float64))
    B = np.abs(conv(b16).reshape(-1)[:k * ldb].reshape(k, ldb)[:, :n].astype(np.float64))
    return T.k_scaled_bound(_tol(op), k, abs(alpha) * (A @ B), 2.0 ** -24, abs(beta) * np.abs(conv(c16).reshape(m, n).astype(np.float64))).reshape(-1)


@pytest.mark.parametrize("m,n,k,lda_pad,ldb_pad", [(128, 256, 64, 0, 0), (128, 256, 256, 0, 0), (256, 512, 192, 0, 0),
                                                   (384, 768, 512, 8, 16), (300, 520, 200, 0, 0), (130, 264, 72, 0, 0)])
def test_gemm_bf16_tensor_route(g, O, m, n, k, lda_pad, ldb_pad):
    lda, ldb = k + lda_pad, n + ldb_pad
    a = _bf16_vals(O, m * lda, 1); b = _bf16_vals(O, k * ldb, 2); c0 = _bf16_vals(O, m * n, 3)
    for alpha, beta in [(1.0, 0.0), (1.5, 0.5)]:
        da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
        g.GEMMBFloat16(False, False, m, n, k, alpha, da, lda, db, ldb, beta, dc, n)
        g.Synchronize()
        got = O.bf16_to_f32(dc.UInt16())
        want = O.bf16_to_f32(O.gemm_bf16(m, n, k, alpha, a, lda, b, ldb, beta, c0.copy(), n))
        nbad, frac_equal = _ulp16_check(got, want, 7, _order_bound(O, a, lda, b, ldb, m, n, k, alpha, beta, c0, O.bf16_to_f32))
        assert nbad == 0, (nbad, frac_equal)
        assert frac_equal > 0.97
        g.Free(da); g.Free(db); g.Free(dc)


def test_gemm_f16_tensor_route(g, O):
    m, n, k = 256, 512, 320
    a = O.f32_to_f16(O.uniform_f32(m * k, 1)); b = O.f32_to_f16(O.uniform_f32(k * n, 2)); c0 = O.f32_to_f16(O.uniform_f32(m * n, 3))
    da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
    g.GEMMFloat16(False, False, m, n, k, 1.5, da, k, db, n, 0.5, dc, n)
    g.Synchronize()
    got = dc.UInt16().view(np.float16).astype(np.float32)
    want = O.gemm_f16(0, 0, m, n, k, 1.5, a, k, b, n, 0.5, c0.copy(), n, threads=8).view(np.float16).astype(np.float32)
    f16 = lambda h: np.ascontiguousarray(h).view(np.float16).astype(np.float32)
    nbad, frac_equal = _ulp16_check(got, want, 10, _order_bound(O, a, k, b, n, m, n, k, 1.5, 0.5, c0, f16, "gemm_f16"), "gemm_f16")
    assert nbad == 0, (nbad, frac_equal)
    assert frac_equal > 0.9
    g.Free(da); g.Free(db); g.Free(dc)


@pytest.mark.parametrize("act", ["relu", "gelu"])
def test_fused_bf16_bias_act_tensor_route(g, O, act):
    m, n, k = 256, 512, 256
    s = 1.0 / np.sqrt(k)
    a = _bf16_vals(O, m * k, 1); b = _bf16_vals(O, k * n, 2, -3 * s, 3 * s); bias = O.uniform_f32(n, 5)
    da, db, dbias, dc = g.to_device(a), g.to_device(b), g.to_device(bias), g.Malloc(m * n * 2)
    if act == "relu":
        g.FusedGEMMBFloat16BiasReLU(m, n, k, 1.0, da, k, db, n, dc, n, dbias)
    else:
        g.FusedGEMMBFloat16BiasGELU(m, n, k, 1.0, da, k, db, n, dc, n, dbias)
    g.Synchronize()
    got = O.bf16_to_f32(dc.UInt16())
    pre = O.bf16_to_f32(a).reshape(m, k).astype(np.float64) @ O.bf16_to_f32(b).reshape(k, n).astype(np.float64)
    v = (pre + bias).astype(np.float32)
    want = np.maximum(v, 0) if act == "relu" else O.gelu(v)
    want = O.bf16_to_f32(O.to_bf16_fast(want)).reshape(-1)
    from guda_b200 import tolerance as T
    assert np.all(np.abs(got - want) <= T.half_ulp_bound(_tol("gemm_bf16_fused"), want, 7, -6))
    for p in (da, db, dbias, dc):
        g.Free(p)


def test_gemm_bf16_large_k_and_modes_agree(g, O):
    """K = 4096 (the C3 depth) on a slab: the tensor route and the exact route agree to 1 bf16 ULP."""
    m, n, k = 256, 512, 4096
    s = 1.0 / np.sqrt(k)
    a = _bf16_vals(O, m * k, 7); b = _bf16_vals(O, k * n, 8, -3 * s, 3 * s)
    da, db, dc1, dc2 = g.to_device(a), g.to_device(b), g.Malloc(m * n * 2), g.Malloc(m * n * 2)
    g.GEMMBFloat16(False, False, m, n, k, 1.0, da, k, db, n, 0.0, dc1, n)
    g.SetGemm16Mode(1)
    try:
        g.GEMMBFloat16(False, False, m, n, k, 1.0, da, k, db, n, 0.0, dc2, n)
    finally:
        g.SetGemm16Mode(0)
    g.Synchronize()
    t = O.bf16_to_f32(dc1.UInt16()); e = O.bf16_to_f32(dc2.UInt16())
    nbad, frac_equal = _ulp16_check(t, e, 7, _order_bound(O, a, k, b, n, m, n, k, 1.0, 0.0, np.zeros(m * n, np.uint16), O.bf16_to_f32))
    assert nbad == 0 and frac_equal > 0.97, (nbad, frac_equal)
    for p in (da, db, dc1, dc2):
        g.Free(p)


@pytest.mark.parametrize("m,n,k", [(128, 128, 32), (256, 384, 512), (300, 260, 1000), (1024, 1024, 1024)])
def test_sgemm_3xtf32_within_stated_bound(g, O, m, n, k):
    a = O.uniform_f32(m * k, 1); b = O.uniform_f32(k * n, 2); c0 = O.uniform_f32(m * n, 3, 0.0, 1.0)
    da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
    g.SetSgemmMode(1)
    try:
        g.GEMM(False, False, m, n, k, 1.5, da, k, db, n, 0.5, dc, n)
        g.Synchronize()
    finally:
        g.SetSgemmMode(0)
    got = dc.Float32().astype(np.float64)
    want = O.sgemm(0, 0, m, n, k, 1.5, a, k, b, n, 0.5, c0.copy(), n, threads=8).astype(np.float64)
    A, B = a.reshape(m, k).astype(np.float64), b.reshape(k, n).astype(np.float64)
    truth = 1.5 * (A @ B) + 0.5 * c0.reshape(m, n)
    from guda_b200 import tolerance as T
    bound = T.k_scaled_bound(_tol("gemm_3xtf32"), k, 1.5 * (np.abs(A) @ np.abs(B)), 2.0 ** -24, 0.5 * np.abs(c0.reshape(m, n)))
    assert np.all(np.abs(got.reshape(m, n) - want.reshape(m, n)) <= bound)
    err_tc = np.max(np.abs(got.reshape(m, n) - truth)); err_ref = np.max(np.abs(want.reshape(m, n) - truth))
    print("3xTF32 max abs err vs fp64: %.3e   reference FMA chain: %.3e" % (err_tc, err_ref))
    assert err_tc <= max(16 * err_ref, 1e-5)
    g.Free(da); g.Free(db); g.Free(dc)
