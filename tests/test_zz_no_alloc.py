"""guda_sdot_device_static: the DOT entry point that allocates nothing (include/guda_b200.h; no counterpart in the reference -- added
for callers that cannot call cudaMalloc).  Checked through the C ABI against the existing entry
point guda_sdot_device (and the oracle), on calls in a row that share the one static scratch array.  Named test_zz_* so that pytest
collects it LAST: a failure here must not hide any other test.  Passes alone in a fresh process (it
initialises the library itself) and after the rest of the suite (it assumes nothing about the scratch array's contents)."""
import numpy as np
import pytest

from tests.helpers import TOL_REDUCE

pytestmark = pytest.mark.gpu

MAX_N = 1 << 28                                         # GUDA_SDOT_STATIC_MAX_N


@pytest.fixture(scope="module")
def g():
    import guda_b200 as G
    G.Init(0)
    return G


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


def _both(g, n, x, incx, y, incy, out, slot, stream=None):
    """The same DOT through the existing entry point (into out[2*slot]) and the static-scratch one (into out[2*slot+1])."""
    L = g.load()
    h = None if stream is None else stream.handle
    g.capi.check(L.guda_sdot_device(n, x.ptr, incx, y.ptr, incy, out.ptr + 8 * slot, h), "SdotDevice")
    g.capi.check(L.guda_sdot_device_static(n, x.ptr, incx, y.ptr, incy, out.ptr + 8 * slot + 4, h), "SdotDeviceStatic")


def test_static_scratch_dot_matches_existing_entry_point_on_calls_in_a_row(g, O):
    # a full-wave call first, then smaller ones: the later calls see the earlier calls' partials and ticket in the shared array
    cases = [((1 << 22) + 3, 1, 1), (1 << 20, 1, 1), (1000, 1, 1), (1, 1, 1), (4099, 2, -3), ((1 << 22) + 3, 1, 1)]
    nmax = max(n * max(abs(ix), abs(iy)) for n, ix, iy in cases)
    xh, yh = O.uniform_f32(nmax, 9101), O.uniform_f32(nmax, 9102)
    x, y = g.to_device(xh), g.to_device(yh)
    out = g.Malloc(8 * len(cases))
    g.Memset(out, 0xFF, 8 * len(cases))                  # NaN patterns: a call that wrote nothing cannot pass
    for slot, (n, ix, iy) in enumerate(cases):           # back to back on the default stream, no synchronisation in between
        _both(g, n, x, ix, y, iy, out, slot)
    g.Synchronize()
    got = out.Float32()[:2 * len(cases)].reshape(len(cases), 2).copy()
    for slot, (n, ix, iy) in enumerate(cases):
        existing, static = np.float32(got[slot, 0]), np.float32(got[slot, 1])
        want = np.float32(O.sdot(n, xh, ix, yh, iy))
        assert np.isfinite(static), (n, ix, iy)
        assert O.near_equal(existing, static, TOL_REDUCE), (n, ix, iy, existing, static)
        assert O.near_equal(want, static, TOL_REDUCE), (n, ix, iy, want, static)
    # same kernel, same grid, same fold order: on one device the two entry points agree to the bit as well
    assert np.array_equal(got[:, 0].view(np.uint32), got[:, 1].view(np.uint32))
    for p in (x, y, out):
        g.Free(p)


def test_static_scratch_dot_allocates_no_device_memory(g, O):
    n = 1 << 21
    x, y = g.to_device(O.uniform_f32(n, 9201)), g.to_device(O.uniform_f32(n, 9202))
    out = g.Malloc(64)
    st = g.CreateStream()
    L = g.load()
    g.capi.check(L.guda_sdot_device_static(n, x.ptr, 1, y.ptr, 1, out.ptr, st.handle), "SdotDeviceStatic")   # module load, symbol lookup
    st.Synchronize()
    free_before, _ = g.MemGetInfo()
    stats_before = g.GetMemoryStats()
    for rep in range(8):
        g.capi.check(L.guda_sdot_device_static(n - rep, x.ptr, 1, y.ptr, 1, out.ptr + 4 * rep, st.handle), "SdotDeviceStatic")
    st.Synchronize()
    free_after, _ = g.MemGetInfo()
    assert free_after == free_before
    assert g.GetMemoryStats() == stats_before
    got = out.Float32()[:8].copy()
    for rep in range(8):
        want = np.float32(g.Sdot(n - rep, x, 1, y, 1, st))
        assert O.near_equal(want, np.float32(got[rep]), TOL_REDUCE), (rep, want, got[rep])
    st.Destroy()
    for p in (x, y, out):
        g.Free(p)


def test_static_scratch_dot_limits_and_edge_cases(g):
    L = g.load()
    out = g.Malloc(16)
    x = g.Malloc(64)
    g.Memset(out, 0xFF, 16)
    # beyond the stated maximum: refused before anything is launched or dereferenced
    rc = L.guda_sdot_device_static(MAX_N + 1, x.ptr, 1, x.ptr, 1, out.ptr, None)
    assert rc == g.capi.ERR_INVALID_ARG
    assert "exceeds the maximum" in L.guda_last_error().decode()
    with pytest.raises(g.GUDAError, match="exceeds the maximum"):
        g.capi.check(L.guda_sdot_device_static(2 ** 31 - 1, x.ptr, 1, x.ptr, 1, out.ptr, None), "SdotDeviceStatic")
    # the reference's argument panics keep their strings
    with pytest.raises(g.capi.BlasPanic, match="zero x index increment"):
        g.capi.check(L.guda_sdot_device_static(4, x.ptr, 0, x.ptr, 1, out.ptr, None), "SdotDeviceStatic")
    with pytest.raises(g.capi.BlasPanic, match="n < 0"):
        g.capi.check(L.guda_sdot_device_static(-1, x.ptr, 1, x.ptr, 1, out.ptr, None), "SdotDeviceStatic")
    with pytest.raises(g.GUDAError, match="null result pointer"):
        g.capi.check(L.guda_sdot_device_static(4, x.ptr, 1, x.ptr, 1, None, None), "SdotDeviceStatic")
    # n == 0: the result is 0 (level1float32_sdot.go:25-30), still without touching x / y
    g.capi.check(L.guda_sdot_device_static(0, None, 1, None, 1, out.ptr, None), "SdotDeviceStatic")
    g.Synchronize()
    assert out.Float32()[0].view(np.uint32) == 0
    g.Free(out); g.Free(x)
