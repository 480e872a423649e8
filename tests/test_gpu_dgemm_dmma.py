"""GPU parity for DGEMM on the FP64 tensor pipe (DMMA m8n8k4, guda_b200/csrc/dgemm_dmma.cu).

The claim under test: chaining m8n8k4 DMMAs over k reproduces the reference's sequential-k FMA chain
(compute/dgemm.go:252-260,321-333; compute/dgemm_fixed.go + compute/asm/f64/gemm_kernel_4x4_amd64.s) BIT FOR BIT,
so the tensor route needs no tolerance.  Every case is compared with the oracle with np.array_equal and with the
SIMT route (SetDgemmMode(1)).
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


def _aligned_copy(c0):
    base = np.zeros(c0.size + 8, dtype=np.float64)
    start = (-base.ctypes.data // 8) % 4              # host C 32-byte aligned, like device memory
    ch = base[start:start + c0.size]
    ch[:] = c0
    return ch


def _gpu_dgemm(g, mode, ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c0, ldc):
    g.SetDgemmMode(mode)
    try:
        da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
        g.Dgemm(bool(ta), bool(tb), m, n, k, alpha, da, lda, db, ldb, beta, dc, ldc)
        g.Synchronize()
        out = dc.Float64()
        assert np.array_equal(da.Float64(), a, equal_nan=True) and np.array_equal(db.Float64(), b, equal_nan=True)
        g.Free(da); g.Free(db); g.Free(dc)
    finally:
        g.SetDgemmMode(0)
    return out


# m, n >= 32 and k >= 4 with even leading dimensions take the DMMA kernel; k % 4 != 0 exercises the scalar tail,
# m % 4 / n % 4 != 0 the unfused-edge split of dgemm_fixed.go, 300 x 260 the partial CTA tiles
SHAPES = [(32, 32, 4), (64, 64, 64), (128, 128, 16), (128, 128, 100), (130, 66, 40), (256, 384, 130), (300, 260, 257),
          (67, 36, 29), (512, 512, 512), (129, 1024, 67)]


@pytest.mark.parametrize("m,n,k", SHAPES)
@pytest.mark.parametrize("ta,alpha,beta", [(0, 1.0, 0.0), (0, 1.0, 0.5), (0, 2.0, 0.5), (1, 2.0, 0.5), (1, 1.0, 1.0), (0, -0.75, 0.0)])
def test_dmma_bit_exact_vs_oracle_and_simt(g, O, m, n, k, ta, alpha, beta):
    lda = (m if ta else k)
    lda += lda & 1                                     # even leading dimension -> DMMA-eligible
    ldb = n + (n & 1)
    rows_a = k if ta else m
    a = O.uniform_f64(rows_a * lda, 11); b = O.uniform_f64(k * ldb, 12); c0 = O.uniform_f64(m * n, 13)
    want = O.dgemm(ta, 0, m, n, k, alpha, a, lda, b, ldb, beta, _aligned_copy(c0), n)
    tensor = _gpu_dgemm(g, 0, ta, 0, m, n, k, alpha, a, lda, b, ldb, beta, c0, n)
    simt = _gpu_dgemm(g, 1, ta, 0, m, n, k, alpha, a, lda, b, ldb, beta, c0, n)
    assert np.array_equal(tensor, want), int(np.sum(tensor != want))
    assert np.array_equal(simt, want), int(np.sum(simt != want))


def test_dmma_zero_skip_semantics(g, O):
    """`if tmp != 0` (compute/dgemm.go:256): a zero in alpha*A must skip an Inf/NaN in B instead of producing NaN, and an
    all-zero A row must leave beta*c (including a negative zero) untouched.  Exercises the in-kernel exact re-computation."""
    m, n, k = 128, 128, 64
    a = O.uniform_f64(m * k, 21); b = O.uniform_f64(k * n, 22); c0 = O.uniform_f64(m * n, 23)
    a = a.reshape(m, k); b = b.reshape(k, n); c0 = c0.reshape(m, n)
    a[:, 5] = 0.0                      # column 5 of A is zero ...
    b[5, 7] = np.inf                   # ... so these never meet
    b[5, 9] = np.nan
    a[3, :] = 0.0                      # row 3: every term skipped
    c0[3, 4] = -0.0
    a, b, c0 = a.ravel().copy(), b.ravel().copy(), c0.ravel().copy()
    for ta, alpha, beta in [(0, 2.0, 1.0), (0, 2.0, 0.5)]:
        want = O.dgemm(ta, 0, m, n, k, alpha, a, k, b, n, beta, _aligned_copy(c0), n)
        assert np.all(np.isfinite(want))
        tensor = _gpu_dgemm(g, 0, ta, 0, m, n, k, alpha, a, k, b, n, beta, c0, n)
        assert np.array_equal(tensor, want)
        assert np.array_equal(np.signbit(tensor), np.signbit(want))
    # alpha == 1 NN runs dgemm_fixed.go's kernels, which have NO zero test: 0 * Inf = NaN there
    want = O.dgemm(0, 0, m, n, k, 1.0, a, k, b, n, 0.0, _aligned_copy(c0), n)
    tensor = _gpu_dgemm(g, 0, 0, 0, m, n, k, 1.0, a, k, b, n, 0.0, c0, n)
    assert np.isnan(want.reshape(m, n)[0, 7]) and np.array_equal(tensor, want, equal_nan=True)


def test_dmma_nan_inf_propagate_like_fma(g, O):
    m, n, k = 64, 96, 32
    a = O.uniform_f64(m * k, 31); b = O.uniform_f64(k * n, 32); c0 = np.zeros(m * n)
    a[17] = np.inf; b[40] = -np.inf; a[300] = np.nan
    want = O.dgemm(0, 0, m, n, k, 1.0, a, k, b, n, 0.0, _aligned_copy(c0), n)
    tensor = _gpu_dgemm(g, 0, 0, 0, m, n, k, 1.0, a, k, b, n, 0.0, c0, n)
    assert np.array_equal(tensor, want, equal_nan=True)


def test_dmma_cancellation_and_subnormals(g, O):
    """Wide dynamic range: products that cancel and subnormal terms round once per FMA, as the chain does."""
    m, n, k = 64, 64, 48
    rng = np.random.default_rng(5)
    a = (rng.standard_normal(m * k) * 10.0 ** rng.integers(-160, 160, m * k)).astype(np.float64)
    b = (rng.standard_normal(k * n) * 10.0 ** rng.integers(-160, 140, k * n)).astype(np.float64)
    c0 = rng.standard_normal(m * n)
    for alpha in (1.0, 3.0):
        want = O.dgemm(0, 0, m, n, k, alpha, a, k, b, n, 1.0, _aligned_copy(c0), n)
        tensor = _gpu_dgemm(g, 0, 0, 0, m, n, k, alpha, a, k, b, n, 1.0, c0, n)
        assert np.array_equal(tensor, want, equal_nan=True)


def test_dmma_sub_views_and_padded_ld(g, O):
    """Operands that are sub-blocks of larger matrices (ld > width), as the 2-D partition uses them."""
    m, n, k, lda, ldb, ldc = 160, 192, 72, 100, 260, 200
    a = O.uniform_f64(m * lda, 41); b = O.uniform_f64(k * ldb, 42); c0 = O.uniform_f64(m * ldc, 43)
    want = O.dgemm(0, 0, m, n, k, 2.5, a, lda, b, ldb, 0.25, _aligned_copy(c0), ldc)
    tensor = _gpu_dgemm(g, 0, 0, 0, m, n, k, 2.5, a, lda, b, ldb, 0.25, c0, ldc)
    assert np.array_equal(tensor, want)


def test_dgemm_2048_sampled_and_modes_identical(g, O):
    """C4 at reduced size: tensor and SIMT routes give identical bits; 512 sampled elements recomputed with math.fma-equivalent
    chains on the host (the oracle on a row/column sub-problem)."""
    n = 2048
    a = O.uniform_f64(n * n, 51); b = O.uniform_f64(n * n, 52)
    c0 = np.zeros(n * n)
    tensor = _gpu_dgemm(g, 0, 0, 0, n, n, n, 1.0, a, n, b, n, 0.0, c0, n)
    simt = _gpu_dgemm(g, 1, 0, 0, n, n, n, 1.0, a, n, b, n, 0.0, c0, n)
    assert np.array_equal(tensor, simt)
    rows = [0, 5, 777, 2047]
    sub_a = np.ascontiguousarray(a.reshape(n, n)[rows, :]).ravel()
    want = O.dgemm(0, 0, len(rows), n, n, 1.0, sub_a, n, b, n, 0.0, _aligned_copy(np.zeros(len(rows) * n)), n)
    assert np.array_equal(tensor.reshape(n, n)[rows, :].ravel(), want)


def test_dgemm_k_blocked_launches_same_bits(g, O):
    """K >= 8192 with a large C is run as 4096-deep K-blocked launches (L2 residency): resuming the chain from the stored fp64
    accumulator must not change a bit.  Oracle on a row subset; tensor vs SIMT on everything."""
    m = n = 2048
    k = 8200
    a = O.uniform_f64(m * k, 61); b = O.uniform_f64(k * n, 62); c0 = O.uniform_f64(m * n, 63)
    tensor = _gpu_dgemm(g, 0, 0, 0, m, n, k, 1.0, a, k, b, n, 0.5, c0, n)
    simt = _gpu_dgemm(g, 1, 0, 0, m, n, k, 1.0, a, k, b, n, 0.5, c0, n)
    assert np.array_equal(tensor, simt)
    rows = [0, 4, 1023, 2044]                      # whole 4-row groups so that dgemm_fixed.go's 4x4 tiling applies
    rows = [r + d for r in rows for d in range(4)]
    sub_a = np.ascontiguousarray(a.reshape(m, k)[rows, :]).ravel()
    sub_c = np.ascontiguousarray(c0.reshape(m, n)[rows, :]).ravel()
    want = O.dgemm(0, 0, len(rows), n, k, 1.0, sub_a, k, b, n, 0.5, _aligned_copy(sub_c), n)
    assert np.array_equal(tensor.reshape(m, n)[rows, :].ravel(), want)
