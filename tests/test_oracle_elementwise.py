"""CPU tests pinning oracle/elementwise.c on the reference's own tables: TestSigmoidAccuracy (activations_test.go:8-44) and
the closed form of TestFusedOperations (guda_test.go:193-240)."""
import os
import subprocess

import numpy as np

from oracle import oracle as O

# activations_test.go:13-23: input, float64 expectation, tolerance
SIGMOID_CASES = [(0.0, 0.5, 1e-6), (1.0, 0.7310585786300049, 1e-5), (-1.0, 0.2689414213699951, 1e-5), (2.0, 0.8807970779778823, 1e-5),
                 (-2.0, 0.11920292202211757, 1e-5), (5.0, 0.9933071490757153, 1e-5), (-5.0, 0.006692850924284856, 1e-5),
                 (10.0, 0.9999546021312976, 1e-4), (-10.0, 4.5397868702434395e-05, 1e-4)]


def test_sigmoid_reference_table():
    for x, want, tol in SIGMOID_CASES:
        got = float(O.ew_sigmoid(np.array([x], np.float32))[0])
        assert abs(got - want) <= tol and abs(got - 1.0 / (1.0 + np.exp(-x))) <= tol


def test_sigmoid_saturation_and_specials():
    x = np.array([-10.000001, 10.000001, -np.inf, np.inf, -0.0, np.nan], np.float32)
    y = O.ew_sigmoid(x)
    assert y[0] == 0 and y[1] == 1 and y[2] == 0 and y[3] == 1 and y[4] == 0.5 and np.isnan(y[5])


def test_fused_operations_closed_form():
    """guda_test.go:193-240: y = ReLU(2*x + bias), N = 1000, tolerance 1e-5 (here: the exact float32 statement sequence)."""
    rng = np.random.default_rng(7)
    x = (rng.random(1000, dtype=np.float32) * 2 - 1).astype(np.float32)
    bias = (rng.random(1000, dtype=np.float32) * np.float32(0.1)).astype(np.float32)
    got = O.fused_execute(x, [bias], [(3, 2.0, 0, 0), (0, 1.0, 0, 0), (4, 0.0, 0, 0)])
    want = np.maximum(np.float32(2.0) * x + np.float32(1.0) * bias, np.float32(0))
    assert np.array_equal(got, want.astype(np.float32))
    assert np.all(np.abs(got - np.maximum(2.0 * x.astype(np.float64) + bias, 0)) <= 1e-5)


def test_fused_broadcast_index_and_operand_skipping():
    x = np.arange(12, dtype=np.float32)
    b = np.array([10, 20, 30], np.float32)
    # shape [4, 3], BroadcastDim 1: broadcastSize = 3, index (idx / 3) % 3 (fused.go:152-161 as written)
    got = O.fused_execute(x, [b], [(0, 1.0, 3, 3)])
    assert np.array_equal(got, x + b[(np.arange(12) // 3) % 3])
    # a second binary op with no operand left is skipped (fused.go:148)
    got = O.fused_execute(x, [np.ones(12, np.float32)], [(0, 2.0, 0, 0), (1, 0.0, 0, 0)])
    assert np.array_equal(got, x + 2)


def test_elementwise_statement_semantics():
    a = np.array([1.5, -0.0, np.inf, 3e38, 1e-45], np.float32)
    b = np.array([2.25, 0.0, -np.inf, 3e38, 0.5], np.float32)
    with np.errstate(all="ignore"):
        assert np.array_equal(O.ew_add(a, b), a + b, equal_nan=True)
        assert np.array_equal(O.ew_mul(a, b), a * b, equal_nan=True)
    r = O.ew_relu(np.array([-1.0, -0.0, 0.0, 2.0, np.nan, -np.inf], np.float32))
    assert r[0] == 0 and np.signbit(r[1]) and r[3] == 2 and np.isnan(r[4]) and r[5] == 0
    # LinearReLU / GELU op lists (fused.go:252-267) against the scalar restatements
    x = O.uniform_f32(1000, 3, -4, 4)
    lin = O.fused_execute(x, [], [(3, 1.7, 0, 0), (2, -0.3, 0, 0), (4, 0, 0, 0)])
    assert np.array_equal(lin, np.maximum(x * np.float32(1.7) + np.float32(-0.3), np.float32(0)))
    assert np.array_equal(O.fused_execute(x, [], [(6, 0, 0, 0)]), O.gelu(x))


def test_device_exp_restatement_is_exhaustively_equivalent(tmp_path):
    """The device ExpFloat32 replaces x / ln2 by x * (1 / ln2) and ldexp by a float multiply; oracle/expcheck.c walks every
    float32 input of the reduction (2.2e9 values, a few seconds on the host cores) and must find no mismatch."""
    src = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "expcheck.c")
    exe = str(tmp_path / "expcheck")
    subprocess.check_call(["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-o", exe, src, "-lm"])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout
    assert "floor mismatches 0  scale mismatches 0" in out.stdout


def test_reduce_axis_is_the_sequential_chain():
    """reduce.go:105-178: Sum from float32(0) in index order; Max from the first element with `val > max`."""
    m, n = 7, 5
    x = O.uniform_f32(m * n, 9, -100, 100)
    X = x.reshape(m, n)
    for axis in (0, 1):
        want = []
        for o in range(n if axis == 0 else m):
            col = X[:, o] if axis == 0 else X[o, :]
            s = np.float32(0)
            for v in col:
                s = np.float32(s + v)
            want.append(s)
        assert np.array_equal(O.reduce_axis(x, m, n, axis), np.array(want, np.float32))
        assert np.array_equal(O.reduce_axis(x, m, n, axis, True), X.max(axis=axis))
    y = np.array([np.nan, 1.0, 2.0, -0.0, 0.0, np.nan], np.float32)          # rows: [nan 1 2], [-0 0 nan]
    r = O.reduce_axis(y, 2, 3, 1, True)
    assert np.isnan(r[0]) and r[1] == 0 and np.signbit(r[1])                  # leading NaN stays; of equal zeros the first wins; later NaN skipped


def test_topk_selection_sort_semantics():
    """reduce.go:247-291: ties come out in the order the swaps leave them; a NaN is picked only when its round starts on it."""
    v, i = O.topk(np.array([3, 1, 3, np.nan, 2, 3, 0], np.float32), 4)
    assert list(i) == [0, 2, 5, 3] and list(v[:3]) == [3, 3, 3] and np.isnan(v[3])
    v, i = O.topk(np.array([np.nan, 5, 1], np.float32), 2)
    assert list(i) == [0, 1] and np.isnan(v[0]) and v[1] == 5
    x = O.uniform_f32(1000, 5)
    v, i = O.topk(x, 10)
    order = np.argsort(-x, kind="stable")[:10]
    assert np.array_equal(i, order) and np.array_equal(v, x[order])
    v, i = O.topk(x[:3], 7)                                  # k clamped to n
    assert len(v) == 3
