"""The TMA-fed exact SGEMM pipeline (guda_b200/csrc/sgemm_pipe.cu) against the oracle's FMA chains (compute/sgemm.go:139-163,
288-325), bit for bit, for every selectable configuration: alpha / beta forms, the TN form, zeros in alpha*A meeting NaN / Inf /
-0 (the `tmp != 0` skip), the K-chunked path (k >= 16384), the fused bias+ReLU / bias+GELU epilogues, the scatter epilogue, and
equality with the register-staged kernel.  Shapes are multiples of the largest tile so that no configuration silently declines."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
CFGS = [1, 2, 3]


@pytest.fixture(scope="module")
def g():
    import guda_b200 as G
    G.Init(0)
    yield G
    G.SetSgemmPipe(-1, 1)


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


def _bits(x):
    return np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)


def _same(got, want):
    """bit-for-bit, except that a NaN matches any NaN (IEEE 754 leaves the payload open: x86 FMA propagates an operand's, the GPU
    returns the canonical quiet NaN) -- the convention of tests/test_gpu_gemm.py's equal_nan comparisons."""
    g_, w_ = np.ascontiguousarray(got, dtype=np.float32).ravel(), np.ascontiguousarray(want, dtype=np.float32).ravel()
    return bool(np.all((g_.view(np.uint32) == w_.view(np.uint32)) | (np.isnan(g_) & np.isnan(w_))))


def _gemm(g, cfg, ta, m, n, k, alpha, a, lda, b, ldb, beta, c0, ldc):
    g.SetSgemmPipe(cfg, 1)
    da, db, dc = g.to_device(a), g.to_device(b), g.to_device(c0)
    g.GEMM(bool(ta), False, m, n, k, alpha, da, lda, db, ldb, beta, dc, ldc)
    g.Synchronize()
    out = dc.Float32().copy()
    for p in (da, db, dc):
        g.Free(p)
    return out


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("m,n,k,ta,alpha,beta", [(128, 128, 32, 0, 1.0, 0.0), (256, 384, 96, 0, 1.5, 0.5), (384, 256, 160, 1, 0.75, 1.0),
                                                 (128, 256, 1056, 0, -2.0, 0.25), (512, 512, 512, 0, 1.0, 0.0), (256, 128, 64, 1, 1.0, 0.0)])
def test_pipe_bit_exact_vs_oracle(g, O, cfg, m, n, k, ta, alpha, beta):
    lda, ldb, ldc = (m if ta else k) + 4, n + 8, n + 4                     # padded leading dimensions (16-byte multiples)
    a = O.uniform_f32((k if ta else m) * lda, 11 + m); b = O.uniform_f32(k * ldb, 12 + n); c0 = O.uniform_f32(m * ldc, 13, 0.0, 1.0)
    got = _gemm(g, cfg, ta, m, n, k, alpha, a, lda, b, ldb, beta, c0, ldc)
    want = O.sgemm(ta, 0, m, n, k, alpha, a, lda, b, ldb, beta, c0.copy(), ldc, threads=8)
    assert np.array_equal(_bits(got), _bits(want))


@pytest.mark.parametrize("cfg", CFGS)
def test_pipe_zero_skip_nan_inf_negative_zero(g, O, cfg):
    """compute/sgemm.go:301 `if tmp != 0`: a zero in alpha*A must not touch a NaN / Inf in B, and must leave a -0 accumulator alone;
    the pre-pass flags the 64 x 16 blocks that hold zeros and only those k-tiles run the predicated loop."""
    m, n, k = 256, 256, 192
    rng = np.random.default_rng(7)
    a = (rng.random((m, k), dtype=np.float32) * 2 - 1)
    a[rng.random((m, k)) < 0.02] = 0.0
    a[64:128, :] = np.where(rng.random((64, k)) < 0.5, 0.0, a[64:128, :])
    a[130, :] = 0.0                                                         # an all-zero row: c stays exactly beta*c (sign of zero included)
    a[200, 5] = 1e-30                                                       # alpha * a underflows to zero only after scaling
    b = (rng.random((k, n), dtype=np.float32) * 2 - 1)
    b[3, 5] = np.inf; b[7, 9] = np.nan; b[100, 200] = -np.inf
    c0 = (rng.random((m, n), dtype=np.float32) * 2 - 1)
    c0[130, :64] = -0.0
    for alpha, beta in ((1.0, 1.0), (1e-20, 1.0), (2.0, 0.0)):
        got = _gemm(g, cfg, 0, m, n, k, alpha, a.ravel(), k, b.ravel(), n, beta, c0.ravel(), n)
        want = O.sgemm(0, 0, m, n, k, alpha, a.ravel(), k, b.ravel(), n, beta, c0.ravel().copy(), n, threads=8)
        assert _same(got, want), (alpha, beta)
        assert 0 < np.isnan(want).sum() < m * n // 8 and np.isinf(want).sum() > 0      # the NaN / Inf stayed in their columns
    assert np.signbit(got.reshape(m, n)[130, 0]) == np.signbit(np.float32(0.0))     # beta == 0 zeroes C first


@pytest.mark.parametrize("cfg", [1, 3])
def test_pipe_k_chunked_chain_resumes_bit_exact(g, O, cfg):
    """k >= 16384 runs as 8192-deep launches resuming the stored chains with beta = 1 (sgemm_pipe.cu / gemm.cu launch_simt): same bits."""
    m, n, k = 128, 128, 16384 + 8192 + 64
    a = O.uniform_f32(m * k, 31); b = O.uniform_f32(k * n, 32); c0 = O.uniform_f32(m * n, 33)
    got = _gemm(g, cfg, 0, m, n, k, 1.0, a, k, b, n, 0.5, c0, n)
    want = O.sgemm(0, 0, m, n, k, 1.0, a, k, b, n, 0.5, c0.copy(), n, threads=8)
    assert np.array_equal(_bits(got), _bits(want))


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("act", ["relu", "gelu"])
def test_pipe_fused_epilogues_bit_exact(g, O, cfg, act):
    m, n, k = 256, 256, 320
    s = 1.0 / np.sqrt(k)
    a = O.uniform_f32(m * k, 41); b = O.uniform_f32(k * n, 42, -3 * s, 3 * s); bias = O.uniform_f32(n, 43)
    g.SetSgemmPipe(cfg, 1)
    da, db, dbias, dc = g.to_device(a), g.to_device(b), g.to_device(bias), g.Malloc(m * n * 4)
    c = np.zeros(m * n, np.float32)
    if act == "relu":
        g.FusedGEMMBiasReLU(False, False, m, n, k, 1.25, da, k, db, n, 0.0, dc, n, dbias)
        want = O.fused_gemm_bias_relu(0, 0, m, n, k, 1.25, a, k, b, n, 0.0, c, n, bias, threads=8)
    else:
        g.FusedGEMMBiasGELU(False, False, m, n, k, 1.25, da, k, db, n, 0.0, dc, n, dbias)
        want = O.fused_gemm_bias_gelu(0, 0, m, n, k, 1.25, a, k, b, n, 0.0, c, n, bias, threads=8)
    g.Synchronize()
    assert np.array_equal(_bits(dc.Float32()), _bits(want))
    for p in (da, db, dbias, dc):
        g.Free(p)


@pytest.mark.parametrize("cfg", [1, 3])
def test_pipe_scatter_epilogue(g, O, cfg):
    """guda_sgemm_scatter through the pipeline: the block lands at (row0, col0) of two larger matrices (on one GPU both are local;
    across GPUs they are peer mappings -- tests/multi_gpu_check.py) and, with c == NULL, nowhere else."""
    m, n, k, ld, row0, col0 = 128, 256, 128, 640, 64, 128
    a = O.uniform_f32(m * k, 51); b = O.uniform_f32(k * n, 52)
    g.SetSgemmPipe(cfg, 1)
    L = g.load()
    da, db = g.to_device(a), g.to_device(b)
    d1, d2 = g.Malloc(320 * ld * 4), g.Malloc(320 * ld * 4)
    g.Memset(d1, 0xFF, 320 * ld * 4); g.Memset(d2, 0xFF, 320 * ld * 4)
    dests = (C.c_void_p * 2)(d1.ptr, d2.ptr)
    g.capi.check(L.guda_sgemm_scatter(0, 0, m, n, k, 1.0, da.ptr, k, db.ptr, n, 0.0, None, n, dests, 2, ld, row0, col0, None), "scatter")
    g.Synchronize()
    want = O.sgemm(0, 0, m, n, k, 1.0, a, k, b, n, 0.0, np.zeros(m * n, np.float32), n, threads=8).reshape(m, n)
    for d in (d1, d2):
        full = d.Float32().reshape(320, ld)
        assert np.array_equal(_bits(full[row0:row0 + m, col0:col0 + n]), _bits(want))
        mask = np.ones((320, ld), bool); mask[row0:row0 + m, col0:col0 + n] = False
        assert np.all(full.view(np.uint32)[mask] == 0xFFFFFFFF)             # nothing outside the block was written
    for p in (da, db, d1, d2):
        g.Free(p)


def test_pipe_equals_register_staged_kernel_and_auto_choice(g, O):
    """Automatic choice (the default) at 1024^3 takes the pipeline; forced off it takes the register-staged kernel: identical bits."""
    n = 1024
    a = O.uniform_f32(n * n, 61); b = O.uniform_f32(n * n, 62); c0 = np.zeros(n * n, np.float32)
    outs = [_gemm(g, cfg, 0, n, n, n, 1.0, a, n, b, n, 0.0, c0, n) for cfg in (-1, 0, 1)]
    assert np.array_equal(_bits(outs[0]), _bits(outs[1])) and np.array_equal(_bits(outs[0]), _bits(outs[2]))
    g.SetSgemmPipe(-1, 1)


def test_pipe_declines_what_it_cannot_take(g, O):
    """Ragged or unaligned calls fall back to the register-staged kernel whatever configuration is forced -- still bit-exact."""
    g.SetSgemmPipe(1, 1)
    for m, n, k, pad in ((130, 128, 64, 0), (128, 136, 64, 0), (128, 128, 40, 0), (128, 128, 64, 1)):
        lda, ldb, ldc = k + pad, n + pad, n + pad
        a = O.uniform_f32(m * lda, 71); b = O.uniform_f32(k * ldb, 72); c0 = O.uniform_f32(m * ldc, 73)
        got = _gemm(g, 1, 0, m, n, k, 1.5, a, lda, b, ldb, 0.5, c0, ldc)
        want = O.sgemm(0, 0, m, n, k, 1.5, a, lda, b, ldb, 0.5, c0.copy(), ldc, threads=8)
        assert np.array_equal(_bits(got), _bits(want))
    g.SetSgemmPipe(-1, 1)
