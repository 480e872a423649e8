"""GPU parity for Conv2D (guda_b200/csrc/conv.cu) through the C ABI: the reference's own conv tests (conv_test.go:8-260) and
random shapes against the oracle's restatement of conv.go -- bit for bit, since both run the same im2col -> Sgemm FMA chain."""
import numpy as np
import pytest

from tests.helpers import golden

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


def _run(g, p, x, w, b):
    cp = g.ConvParams(**p)
    oh, ow = cp.OutputHeight(), cp.OutputWidth()
    n_out = p["BatchSize"] * p["OutChannels"] * oh * ow
    dx, dw, dout = g.to_device(x), g.to_device(w), g.Malloc(n_out * 4)
    g.Memset(dout, 0xFF, n_out * 4)
    db = g.to_device(b) if b is not None else g.DevicePtr()
    g.Conv2D(dx, dw, db, dout, cp)
    out = dout.Float32()
    assert np.array_equal(dx.Float32(), x) and np.array_equal(dw.Float32(), w)
    for d in (dx, dw, dout) + ((db,) if b is not None else ()):
        g.Free(d)
    return out


@pytest.mark.parametrize("idx", range(4))
def test_reference_conv_cases_on_gpu(g, idx, route):
    case = golden("conv_cases.json")["cases"][idx]
    bias = np.array(case["bias"], dtype=np.float32) if case["bias"] else None
    out = _run(g, case["params"], np.array(case["input"], np.float32), np.array(case["kernel"], np.float32), bias)
    exp = np.array(case["expected"], dtype=np.float32)
    assert np.all(np.abs(out[:exp.size] - exp) <= 1e-5)


def _params(B, C_, H, W, OC, KH, KW, s=1, p=0, d=1, bias=True):
    return {"BatchSize": B, "InChannels": C_, "InHeight": H, "InWidth": W, "OutChannels": OC, "KernelHeight": KH, "KernelWidth": KW,
            "StrideH": s, "StrideW": s, "PadH": p, "PadW": p, "DilationH": d, "DilationW": d, "UseBias": bias}


# conv_test.go:262-300 benchmark shapes (32x32x32 3x3, 64x56x56 3x3 -> 64, 3x224x224 7x7 s2) + ragged / dilated / batched ones
SHAPES = [_params(1, 32, 32, 32, 64, 3, 3, 1, 1), _params(1, 64, 56, 56, 64, 3, 3, 1, 1), _params(1, 3, 224, 224, 64, 7, 7, 2, 3),
          _params(1, 3, 9, 11, 5, 3, 3, 1, 1, bias=False), _params(2, 4, 8, 8, 6, 3, 3, 2, 1), _params(3, 2, 10, 7, 4, 2, 3, 1, 0, 2),
          _params(4, 16, 14, 14, 32, 1, 1), _params(1, 1, 5, 5, 1, 5, 5),
          # tile-edge cases of the implicit-GEMM route: OC over one 64-row tile, N = B*P over / not a multiple of 128 with the
          # batch boundary inside a 4-wide store group (P = 35), K not a multiple of 16, the Sgemm 16^3 fast-path shape (OC = N = K = 16)
          _params(3, 5, 9, 7, 70, 3, 3, 1, 1), _params(5, 3, 7, 5, 130, 3, 2, 1, 1, bias=False), _params(2, 8, 33, 17, 64, 3, 3, 2, 2, 2),
          _params(1, 16, 4, 4, 16, 1, 1), _params(2, 128, 12, 12, 128, 3, 3, 1, 1)]


@pytest.fixture(params=[0, 1, 2, 3, 4], ids=["implicit", "im2col", "implicit-64x128", "implicit-64x64", "implicit-128x128"])
def route(request, g):
    g.SetConvRoute(request.param)
    yield request.param
    g.SetConvRoute(-1)


def test_conv2d_routes_bit_identical_large(g, O):
    """Both routes at a size the oracle does not replay: 64x56x56 -> 64, 3x3, batch 8 (N = 25088), bit for bit."""
    p = _params(8, 64, 56, 56, 64, 3, 3, 1, 1)
    x = O.uniform_f32(8 * 64 * 56 * 56, 11)
    w = O.uniform_f32(64 * 64 * 9, 12)
    w[::5] = 0.0
    b = O.uniform_f32(64, 13)
    outs = []
    for r in (0, 1, 2, 3, 4):
        g.SetConvRoute(r)
        outs.append(_run(g, p, x, w, b))
    g.SetConvRoute(-1)
    for o in outs[1:]:
        assert np.array_equal(outs[0], o)
    assert np.all(np.isfinite(outs[0])) and float(np.abs(outs[0]).max()) > 0


@pytest.mark.parametrize("p", SHAPES)
def test_conv2d_bit_exact_vs_oracle(g, O, p, route):
    x = O.uniform_f32(p["BatchSize"] * p["InChannels"] * p["InHeight"] * p["InWidth"], 1)
    w = O.uniform_f32(p["OutChannels"] * p["InChannels"] * p["KernelHeight"] * p["KernelWidth"], 2)
    w[::7] = 0.0                                    # the Sgemm zero test acts on the weights
    b = O.uniform_f32(p["OutChannels"], 3) if p["UseBias"] else None
    got = _run(g, p, x, w, b)
    want = O.conv2d(x, w, b, p, threads=8)
    assert np.array_equal(got, want)
    direct = O.conv2d(x, w, b, p, direct=True)
    assert np.max(np.abs(got - direct)) <= 2e-5 * max(1.0, float(np.max(np.abs(direct))))


def test_conv2d_validation_errors(g):
    p = _params(1, 1, 4, 4, 1, 3, 3)
    z = g.Malloc(64)
    for field, val in (("StrideH", 0), ("DilationW", 0), ("PadH", -1), ("OutChannels", 0), ("BatchSize", 0)):
        bad = dict(p); bad[field] = val
        with pytest.raises(g.GUDAError) as e:
            g.Conv2D(z, z, None, z, g.ConvParams(**bad))
        assert e.value.type == 2 and "invalid" in e.value.message           # InvalidArgument, conv.go:34-51
    g.Free(z)


@pytest.mark.parametrize("p", SHAPES)
def test_conv2d_direct_bit_exact_vs_oracle(g, O, p):
    """conv.go:231-300 Conv2DDirect: unfused sum over the in-bounds taps, no zero test -- bit for bit, non-finite weights included."""
    x = O.uniform_f32(p["BatchSize"] * p["InChannels"] * p["InHeight"] * p["InWidth"], 21)
    w = O.uniform_f32(p["OutChannels"] * p["InChannels"] * p["KernelHeight"] * p["KernelWidth"], 22)
    w[::7] = 0.0
    if w.size > 3:
        w[3] = np.inf                                # an Inf weight on a padded tap must not turn into 0 * Inf = NaN
    b = O.uniform_f32(p["OutChannels"], 23) if p["UseBias"] else None
    cp = g.ConvParams(**p)
    n_out = p["BatchSize"] * p["OutChannels"] * cp.OutputHeight() * cp.OutputWidth()
    dx, dw, dout = g.to_device(x), g.to_device(w), g.Malloc(n_out * 4)
    g.Memset(dout, 0xFF, n_out * 4)
    db = g.to_device(b) if b is not None else g.DevicePtr()
    g.Conv2DDirect(dx, dw, db, dout, cp)
    got = dout.Float32()[:n_out]
    with np.errstate(all="ignore"):
        want = O.conv2d(x, w, b, p, direct=True)
    nan = np.isnan(want)
    assert np.array_equal(nan, np.isnan(got)) and np.array_equal(got[~nan].view(np.uint32), want[~nan].view(np.uint32))
    for d in (dx, dw, dout) + ((db,) if b is not None else ()):
        g.Free(d)
