"""The architecture-aware tolerance framework: guda_b200/tolerance.py + the C ABI behind it
(guda_b200/csrc/tolerance.cu), pinned on the values and cases the reference holds in tolerance.go:28-54,
tolerance_arch.go:41-127, numerical_parity.go:103-141, tolerance_test.go:8-240,345-400 and tolerance_arch_test.go:27-57.

Five expectations of the reference's own tests are NOT met by the reference's own code, and the restatement follows the code:
  tolerance_test.go:38-44  "Within_RelTol"  1000.0 vs 1000.01 under DefaultTolerance: |a - b| in float32 is 0.010009766, the
                           bound 1000.01 * 1e-5 is 0.0100001 and the values are 164 ULP apart -> false (the test wants true);
  tolerance_test.go:84-90  "Mixed_Inf"      +Inf vs -Inf: diff = +Inf <= larger * RelTol = +Inf -> true (the test wants false);
  tolerance_test.go:99-105 "Outside_ULP"    1.0 vs 1.0 + 5 ULP: 5 ULP at 1.0 is 6e-7, inside RelTol 1e-5 -> true (the test wants false);
  tolerance_test.go:180-186 "Within_Tolerance"  float32(1.00001) - 1 = 1.0013580e-05 > 1.00001 * float32(1e-5) = 1.0000097e-05 and
                           84 ULP -> element 0 is an error, IsAcceptable false (the test wants a pass);
  tolerance_test.go:194-200 "Different_Lengths"  NumErrors = len(expected) but every Max*Error stays 0, so IsAcceptable's second
                           clause (tolerance.go:177-182) is true (the test wants a failure).
The CPU tests never touch a device; the GPU test checks guda_verify_f32 (VerifyFloat32Array on device arrays) against the oracle."""
import math

import numpy as np
import pytest

import guda_b200 as G
from guda_b200 import tolerance as T


def _f32(x):
    return float(np.float32(x))


def test_presets_tolerance_go_28_54():
    d, s, r = G.DefaultTolerance(), G.StrictTolerance(), G.RelaxedTolerance()
    assert (d.AbsTol, d.RelTol, d.ULPTol, d.CheckNaN, d.CheckInf) == (_f32(1e-7), _f32(1e-5), 4, True, True)
    assert (s.AbsTol, s.RelTol, s.ULPTol) == (_f32(1e-9), _f32(1e-7), 1)
    assert (r.AbsTol, r.RelTol, r.ULPTol) == (_f32(1e-5), _f32(1e-3), 16)
    # tolerance_test.go:345-400 TestTolerancePresets ranges
    for tol, amin, amax, rmin, rmax, umin, umax in ((d, 1e-8, 1e-6, 1e-6, 1e-4, 2, 8), (s, 1e-10, 1e-8, 1e-8, 1e-6, 1, 2),
                                                   (r, 1e-6, 1e-4, 1e-4, 1e-2, 8, 32)):
        assert amin <= tol.AbsTol <= amax and rmin <= tol.RelTol <= rmax and umin <= tol.ULPTol <= umax


NEXT = lambda x, k: float(np.uint32(np.float32(x).view(np.uint32) + k).view(np.float32))


@pytest.mark.parametrize("name,a,b,tol,want", [
    ("Exact_Equal", 1.0, 1.0, "default", True), ("Within_AbsTol", 1e-8, 2e-8, "default", True),
    ("Outside_AbsTol", 1e-6, 2e-6, "default", False), ("Both_Zero", 0.0, -0.0, "default", True),
    ("Both_NaN", math.nan, math.nan, "default", True), ("NaN_Not_Checked", math.nan, math.nan, "none", False),
    ("Both_PosInf", math.inf, math.inf, "default", True), ("Both_NegInf", -math.inf, -math.inf, "default", True),
    ("Within_ULP", 1.0, NEXT(1.0, 2), "default", True), ("Outside_ULP_as_coded", 1.0, NEXT(1.0, 5), "default", True),
    ("Outside_every_bound", 1.0, NEXT(1.0, 200), "default", False),
    # the cases where the reference's code and its test disagree (see the module docstring): the code's answer
    ("Within_RelTol_as_coded", 1000.0, 1000.01, "default", False), ("Mixed_Inf_as_coded", math.inf, -math.inf, "default", True),
    ("Within_RelTol_f32_exact", 1000.0, 1000.005, "default", True),
])
def test_float32_near_equal_cases_tolerance_test_go_8_120(name, a, b, tol, want):
    t = G.DefaultTolerance() if tol == "default" else G.ToleranceConfig(CheckNaN=False)
    assert G.Float32NearEqual(a, b, t) is want


def test_float32_ulp_diff_tolerance_test_go_122_163():
    assert G.Float32ULPDiff(1.0, 1.0) == 0
    assert G.Float32ULPDiff(1.0, NEXT(1.0, 1)) == 1
    assert G.Float32ULPDiff(1.0, NEXT(1.0, 2)) == 2
    assert G.Float32ULPDiff(1.0, -1.0) == 2 ** 31 - 1


@pytest.mark.parametrize("name,exp,act,tol,want_pass", [
    ("All_Match", [1, 2, 3, 4], [1, 2, 3, 4], "default", True),
    ("Within_Tolerance_as_coded", [1, 2, 3, 4], [1.00001, 2.00001, 3.00001, 4.00001], "default", False),
    ("Within_Tolerance_half", [1, 2, 3, 4], [1.000005, 2.000005, 3.000005, 4.000005], "default", True),
    ("Outside_Tolerance", [1, 2, 3, 4], [1.1, 2, 3, 4], "default", False),
    ("Different_Lengths", [1, 2, 3], [1, 2], "default", True),            # as coded, see the module docstring
    ("With_NaN", [1, math.nan, 3], [1, math.nan, 3], "default", True),
    ("Accumulated_Error", [1000.0], [1001.0], "relaxed", True),
])
def test_verify_float32_array_tolerance_test_go_165_238(name, exp, act, tol, want_pass):
    t = G.DefaultTolerance() if tol == "default" else G.RelaxedTolerance()
    res = G.VerifyFloat32Array(exp, act, t)
    assert res.IsAcceptable(t) is want_pass, res.String()
    if name == "All_Match":
        assert res.NumErrors == 0 and res.String().startswith("PASS")
    if name == "Different_Lengths":
        assert res.NumErrors == len(exp)
    if name == "Outside_Tolerance":
        assert res.FirstError == 0 and res.NumErrors == 1 and res.String().startswith("FAIL: 1/4")


def test_merge_rule_tolerance_arch_test_go_27_57():
    base = G.ToleranceConfig(AbsTol=1e-7, RelTol=1e-6, ULPTol=2, CheckNaN=True, CheckInf=False)
    merged = G.mergeTolerances(base, G.ToleranceConfig(RelTol=1e-4, ULPTol=16))
    assert merged.AbsTol == _f32(1e-7) and merged.RelTol == _f32(1e-4) and merged.ULPTol == 16
    assert merged.CheckNaN is True and merged.CheckInf is False          # the flags always come from the base
    # the sm_100a fields obey the same non-zero rule
    m2 = G.mergeTolerances(G.ToleranceConfig(AbsTol=1e-6, KScale=2.0, Abs64=1e-14), G.ToleranceConfig(KScale=8.0, HalfULPTol=1))
    assert (m2.AbsTol, m2.KScale, m2.HalfULPTol, m2.Abs64) == (_f32(1e-6), 8.0, 1, 1e-14)


def test_operation_tolerances_per_architecture_tolerance_arch_go_59_127():
    want = {  # operation -> {arch: (abs, rel, ulp)}
        "gemm": {"amd64": (1e-6, 1e-5, 4), "arm64": (1e-5, 1e-4, 16), "generic": (1e-4, 1e-3, 32)},
        "conv2d": {"amd64": (1e-6, 1e-5, 8), "arm64": (1e-5, 1e-4, 32), "generic": (1e-6, 1e-5, 8)},
        "reduce_sum": {"amd64": (1e-5, 1e-4, 16), "arm64": (1e-4, 1e-3, 64), "generic": (1e-5, 1e-4, 16)},
        "softmax": {"amd64": (1e-5, 1e-4, 16), "arm64": (1e-4, 1e-3, 64), "generic": (1e-5, 1e-4, 16)},
        "something_else": {"amd64": (1e-7, 1e-5, 4), "arm64": (1e-7, 1e-5, 4), "generic": (1e-7, 1e-5, 4)},   # default: DefaultTolerance()
    }
    for op, per_arch in want.items():
        for arch, (a, r, u) in per_arch.items():
            t = G.GetOperationTolerance(op, arch)
            assert (t.AbsTol, t.RelTol, t.ULPTol) == (_f32(a), _f32(r), u), (op, arch)
            assert t.CheckNaN and t.CheckInf and t.KScale == 0 and t.HalfULPTol == 0 and t.Abs64 == 0
    with pytest.raises(G.GUDAError):
        G.GetOperationTolerance("gemm", "riscv")
    # the arch-config views agree with GetArchTolerance + the merge rule
    cfg = G.GEMMArchTolerance
    assert cfg.AMD64 is None and cfg.ARM64.ULPTol == 16 and cfg.Generic.ULPTol == 32
    assert G.GetArchTolerance(cfg, "arm64") == G.GetOperationTolerance("gemm", "arm64")
    assert G.GetArchTolerance(cfg, "amd64") == cfg.Base
    assert G.GetArchTolerance(cfg) == G.GetOperationTolerance("gemm")
    assert G.IsARM64() is False


def test_sm_100a_entry():
    """The new architecture entry: the default arch of every lookup, base values kept, extra bounds added."""
    g = G.GetOperationTolerance("gemm")
    assert (g.AbsTol, g.RelTol, g.ULPTol, g.KScale) == (_f32(1e-6), _f32(1e-5), 4, 2.0)
    assert G.GetOperationTolerance("gemm_3xtf32").KScale == 8.0
    for op, half in (("gemm_bf16", 1), ("gemm_f16", 1), ("gemm_bf16_fused", 2)):
        t = G.GetOperationTolerance(op)
        assert t.HalfULPTol == half and t.KScale == 2.0
    d = G.GetOperationTolerance("dgemm")
    assert (d.Abs64, d.Rel64, d.KScale) == (1e-14, 1e-14, 2.0)
    r = G.GetOperationTolerance("dot")
    assert (r.AbsTol, r.RelTol, r.ULPTol) == (_f32(1e-5), _f32(1e-4), 16) and r == G.GetOperationTolerance("reduce_sum")
    assert G.GetOperationTolerance("gemm", "sm_100a") == g
    b = T.k_scaled_bound(g, 1024, np.array([3.0]))
    assert b[0] == 2.0 * 1024 * 2.0 ** -24 * 3.0
    assert T.half_ulp_bound(G.GetOperationTolerance("gemm_bf16"), np.array([1.5]), 7, -14)[0] == 1.5 * 2.0 ** -7


def test_standard_tolerances_numerical_parity_go_132_141():
    st = G.StandardTolerances
    assert (st["add"].AbsTol, st["add"].RelTol, st["add"].ULPTol) == (0, 0, 0) and st["multiply"].ULPTol == 0
    for op in ("gemm", "conv2d"):
        assert (st[op].AbsTol, st[op].RelTol, st[op].ULPTol) == (_f32(1e-6), _f32(1e-6), 2) and st[op].Name == op
    for op in ("reduce_sum", "softmax"):
        assert (st[op].AbsTol, st[op].RelTol, st[op].ULPTol) == (_f32(1e-5), _f32(1e-5), 10)
    for op in ("gelu", "exp"):
        assert (st[op].AbsTol, st[op].RelTol, st[op].ULPTol) == (_f32(1e-4), _f32(1e-4), 100)
    assert len(st) == 8


def test_host_mirror_agrees_with_the_oracle_restatement():
    """guda_b200.tolerance (host mirror) vs oracle/fused.c (the checker's restatement of tolerance.go) on 20k random pairs."""
    from oracle import oracle as O
    rng = np.random.default_rng(3)
    a = (rng.standard_normal(20000) * 10.0 ** rng.integers(-8, 4, 20000)).astype(np.float32)
    b = a.copy()
    k = rng.integers(0, 6, 20000)
    b[k == 1] = (b[k == 1].view(np.uint32) + rng.integers(1, 40, int((k == 1).sum())).astype(np.uint32)).view(np.float32)
    b[k == 2] *= np.float32(1 + 3e-5)
    b[k == 3] += np.float32(5e-6)
    b[k == 4] = -b[k == 4]
    a[:4] = [np.nan, np.inf, -np.inf, np.inf]; b[:4] = [np.nan, np.inf, -np.inf, -np.inf]
    for tol, od in ((G.DefaultTolerance(), {"abs": 1e-7, "rel": 1e-5, "ulp": 4}), (G.GetOperationTolerance("reduce_sum"), {"abs": 1e-5, "rel": 1e-4, "ulp": 16})):
        mine = np.array([G.Float32NearEqual(x, y, tol) for x, y in zip(a[:3000], b[:3000])])
        theirs = np.array([O.near_equal(float(x), float(y), od) for x, y in zip(a[:3000], b[:3000])])
        assert np.array_equal(mine, theirs)
        v = G.VerifyFloat32Array(a[:3000], b[:3000], tol); w = O.verify_f32(a[:3000], b[:3000], od)
        assert (v.NumErrors, v.FirstError, v.MaxULPError, np.float32(v.MaxAbsError), np.float32(v.MaxRelError)) == \
               (w["num_errors"], w["first_error"], w["max_ulp"], np.float32(w["max_abs"]), np.float32(w["max_rel"]))


@pytest.mark.gpu
@pytest.mark.parametrize("n,offset", [(1, 0), (7, 0), (4096, 0), (100003, 0), (100003, 1), (1 << 22, 0)])
def test_verify_f32_on_device_matches_the_oracle(n, offset):
    """guda_verify_f32 = VerifyFloat32Array over two device arrays: every integer and both float maxima are the oracle's bits."""
    from oracle import oracle as O
    G.Init(0)
    rng = np.random.default_rng(n + offset)
    e = (rng.standard_normal(n + offset) * 10.0 ** rng.integers(-6, 3, n + offset)).astype(np.float32)
    a = e.copy()
    k = rng.integers(0, 40, n + offset)
    sel = k == 1
    a[sel] = (a[sel].view(np.uint32) + rng.integers(1, 64, int(sel.sum())).astype(np.uint32)).view(np.float32)
    a[k == 2] *= np.float32(1 + 2e-4)
    a[k == 3] = -a[k == 3]
    if n + offset > 16:
        e[5], a[5] = np.nan, np.nan; e[6], a[6] = np.inf, -np.inf; e[7], a[7] = np.nan, 1.0; e[8], a[8] = 0.0, 1e-3; e[9], a[9] = 1.0, np.inf
    de, da = G.to_device(e), G.to_device(a)
    for tol, od in ((G.DefaultTolerance(), {"abs": 1e-7, "rel": 1e-5, "ulp": 4}), (G.GetOperationTolerance("gemm"), {"abs": 1e-6, "rel": 1e-5, "ulp": 4}),
                    (G.StrictTolerance(), {"abs": 1e-9, "rel": 1e-7, "ulp": 1})):
        got = G.VerifyFloat32Array(G.DevicePtr(de.ptr + 4 * offset, 4 * n), G.DevicePtr(da.ptr + 4 * offset, 4 * n), tol, n=n)
        want = O.verify_f32(e[offset:], a[offset:], od)
        assert (got.NumErrors, got.TotalItems, got.FirstError, got.MaxULPError) == (want["num_errors"], n, want["first_error"], want["max_ulp"])
        assert np.float32(got.MaxAbsError).view(np.uint32) == np.float32(want["max_abs"]).view(np.uint32)
        assert np.float32(got.MaxRelError).view(np.uint32) == np.float32(want["max_rel"]).view(np.uint32)
    G.Free(de); G.Free(da)
