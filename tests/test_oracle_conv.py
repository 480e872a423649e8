"""Oracle Conv2D (oracle/conv.c restating conv.go) against the reference's own conv tests (tests/golden/conv_cases.json,
extracted from conv_test.go:8-260) and against the reference's direct form (conv.go:231-300)."""
import numpy as np
import pytest

from tests.helpers import golden


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    return oracle


@pytest.mark.parametrize("idx", range(4))
def test_reference_conv_cases(O, idx):
    case = golden("conv_cases.json")["cases"][idx]
    bias = np.array(case["bias"], dtype=np.float32) if case["bias"] else None
    for direct in (False, True):
        out = O.conv2d(np.array(case["input"], np.float32), np.array(case["kernel"], np.float32), bias, case["params"], direct=direct)
        exp = np.array(case["expected"], dtype=np.float32)
        assert np.all(np.abs(out[:exp.size] - exp) <= 1e-5), (case["name"], direct)      # conv_test.go tolerance


def _params(B, C_, H, W, OC, KH, KW, s=1, p=0, d=1, bias=True):
    return {"BatchSize": B, "InChannels": C_, "InHeight": H, "InWidth": W, "OutChannels": OC, "KernelHeight": KH, "KernelWidth": KW,
            "StrideH": s, "StrideW": s, "PadH": p, "PadW": p, "DilationH": d, "DilationW": d, "UseBias": bias}


@pytest.mark.parametrize("p", [_params(1, 3, 9, 11, 5, 3, 3, 1, 1), _params(2, 4, 8, 8, 6, 3, 3, 2, 1), _params(3, 2, 10, 7, 4, 2, 3, 1, 0, 2),
                               _params(1, 8, 6, 6, 16, 1, 1, bias=False)])
def test_im2col_route_matches_direct_route(O, p):
    """Both reference routes compute the same convolution; they differ only in fp32 rounding (FMA chain vs unfused sum)."""
    oh, ow = O.conv_out_hw(p)
    x = O.uniform_f32(p["BatchSize"] * p["InChannels"] * p["InHeight"] * p["InWidth"], 1)
    w = O.uniform_f32(p["OutChannels"] * p["InChannels"] * p["KernelHeight"] * p["KernelWidth"], 2)
    b = O.uniform_f32(p["OutChannels"], 3)
    a = O.conv2d(x, w, b, p)
    d = O.conv2d(x, w, b, p, direct=True)
    assert a.size == p["BatchSize"] * p["OutChannels"] * oh * ow
    assert np.max(np.abs(a - d)) <= 1e-5 * max(1.0, float(np.max(np.abs(d))))


def test_invalid_params_are_rejected(O):
    p = _params(1, 1, 4, 4, 1, 3, 3)
    p["StrideH"] = 0
    with pytest.raises(O.OraclePanic):
        O.conv2d(np.zeros(16, np.float32), np.zeros(9, np.float32), None, p)
