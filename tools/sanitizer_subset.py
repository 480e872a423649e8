"""Small-shape pass over the kernels with hand-rolled synchronisation, for compute-sanitizer (memcheck / racecheck):
dot_kernel + reduce_kernel (ticket fold), sgemv_t_kernel + dgemv ring (cp.async + mbarrier), gemm_tc_pair_kernel (tcgen05 CTA pair),
sgemm_prepass_kernel + sgemm_pipe_kernel (TMA ring, rotating producer), dgemm_dmma_kernel (cp.async ring), verify_f32_kernel.
Every result is also checked against the oracle, so a sanitizer-clean run is a correct run.

  compute-sanitizer --tool memcheck  python tools/sanitizer_subset.py
  compute-sanitizer --tool racecheck python tools/sanitizer_subset.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
from oracle import oracle as O
g.Init(0)
ok = True

def check(name, cond):
    global ok
    ok = ok and bool(cond)
    print("%-28s %s" % (name, "ok" if cond else "MISMATCH"), flush=True)

n = 70001
x, y = O.uniform_f32(n, 1), O.uniform_f32(n, 2)
dx, dy = g.to_device(x), g.to_device(y)
want = float(O.sdot(n, x, 1, y, 1))
check("dot_kernel", abs(float(g.DOT(dx, dy, n)) - want) <= 1e-4 * max(1.0, abs(want)))
check("reduce_kernel max/argmax", dx.Max(n) == O.smax(x) and dx.ArgMax(n) == O.sargmax(x))
g.AXPY(2.5, dx, dy, n)
check("axpy_unit_kernel", np.array_equal(dy.Float32(), O.saxpy(n, 2.5, x, 1, y.copy(), 1)))
m, nn = 300, 520
a, vx, vy = O.uniform_f32(m * nn, 3), O.uniform_f32(m, 4), O.uniform_f32(nn, 5)
da, dvx, dvy = g.to_device(a), g.to_device(vx), g.to_device(vy)
g.Sgemv(True, m, nn, 1.5, da, nn, dvx, 1, 0.5, dvy, 1)
check("sgemv_t_kernel", np.array_equal(dvy.Float32(), O.sgemv(1, m, nn, 1.5, a, nn, vx, 1, 0.5, vy.copy(), 1)))
# exact SGEMM: register-staged kernel and the TMA pipeline (pre-pass + ring), zeros in A
mm, nk, kk = 128, 256, 96
a2, b2 = O.uniform_f32(mm * kk, 6), O.uniform_f32(kk * nk, 7)
a2[::97] = 0.0
da2, db2, dc2 = g.to_device(a2), g.to_device(b2), g.Malloc(mm * nk * 4)
wantc = O.sgemm(0, 0, mm, nk, kk, 1.5, a2, kk, b2, nk, 0.0, np.zeros(mm * nk, np.float32), nk)
for cfg in (0, 1, 3):
    g.SetSgemmPipe(cfg, 1)
    g.GEMM(False, False, mm, nk, kk, 1.5, da2, kk, db2, nk, 0.0, dc2, nk)
    g.Synchronize()
    check("sgemm cfg %d" % cfg, np.array_equal(dc2.Float32(), wantc))
g.SetSgemmPipe(-1, 1)
# DGEMM on the FP64 tensor pipe
ad, bd = O.uniform_f64(128 * 64, 8), O.uniform_f64(64 * 128, 9)
dad, dbd, dcd = g.to_device(ad), g.to_device(bd), g.Malloc(128 * 128 * 8)
g.Dgemm(False, False, 128, 128, 64, 1.0, dad, 64, dbd, 128, 0.0, dcd, 128)
g.Synchronize()
check("dgemm_dmma_kernel", np.array_equal(dcd.Float64(), O.dgemm(0, 0, 128, 128, 64, 1.0, ad, 64, bd, 128, 0.0, np.zeros(128 * 128), 128)))
# bf16 GEMM on tcgen05 (CTA pair route needs m, n >= 256)
M, N, K = 256, 256, 64
a16, b16 = O.to_bf16_fast(O.uniform_f32(M * K, 10)), O.to_bf16_fast(O.uniform_f32(K * N, 11))
da16, db16, dc16 = g.to_device(a16), g.to_device(b16), g.Malloc(M * N * 2)
g.GEMMBFloat16(False, False, M, N, K, 1.0, da16, K, db16, N, 0.0, dc16, N)
g.Synchronize()
got = O.bf16_to_f32(dc16.UInt16()); wantb = O.bf16_to_f32(O.gemm_bf16(M, N, K, 1.0, a16, K, b16, N, 0.0, np.zeros(M * N, np.uint16), N))
check("gemm_tc_pair_kernel", np.all(np.abs(got - wantb) <= np.maximum(np.abs(wantb), 2.0 ** -14) * 2.0 ** -7 * 1.01 + 1e-5))
# device verification
e = O.uniform_f32(10001, 12); act = e.copy(); act[::50] *= np.float32(1.001)
de, dact = g.to_device(e), g.to_device(act)
r = g.VerifyFloat32Array(de, dact, g.DefaultTolerance(), n=10001)
w = O.verify_f32(e, act, {"abs": 1e-7, "rel": 1e-5, "ulp": 4})
check("verify_f32_kernel", (r.NumErrors, r.FirstError, r.MaxULPError) == (w["num_errors"], w["first_error"], w["max_ulp"]))
g.Synchronize()
print("sanitizer subset:", "all checks passed" if ok else "FAILED")
sys.exit(0 if ok else 1)
