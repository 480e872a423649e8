(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/r1y_pytest.log 2>&1
cat gpurun_out/r1y_pytest.log
python - <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import guda_b200 as g
g.Init(0)
e0, e1 = g.Event(), g.Event()
n = 1 << 26
a, b, c = g.Malloc(n*4), g.Malloc(n*4), g.Malloc(n*4)
g.FillUniform(a, 1, n, n, 1); g.FillUniform(b, 1, n, n, 2)
def t(fn, it=20):
    for _ in range(3): fn()
    e0.Record()
    for _ in range(it): fn()
    e1.Record(); e1.Synchronize()
    return g.ElapsedTime(e0, e1)/it
fk = g.NewFusedKernel().MulScalar(1.01).Add(0.5).ReLU()
for name, fn, bytes_ in (("Add", lambda: g.Add(a, b, c, n), 12), ("Multiply", lambda: g.Multiply(a, b, c, n), 12), ("ReLU", lambda: g.ReLU(a, n), 8),
                         ("Sigmoid", lambda: g.Sigmoid(c, n), 8), ("GELU", lambda: g.GELU(c, n), 8), ("LinearReLU", lambda: g.LinearReLU(c, 1.01, 0.1, n), 8),
                         ("Fused mul+add+relu", lambda: fk.Execute(c, [b], n), 12)):
    ms = t(fn)
    print("%s n=2^26: %.3f ms %.0f GB/s" % (name, ms, bytes_*n/ms/1e6), flush=True)
M, K, N = 8192, 4096, 16384
A, B, Cc, bias = g.Malloc(M*K*4), g.Malloc(K*N*4), g.Malloc(M*N*4), g.Malloc(N*4)
g.FillUniform(A, M, K, K, 5); g.FillUniform(B, K, N, N, 6, lo=-0.05, hi=0.05); g.FillUniform(bias, 1, N, N, 7)
ms = t(lambda: g.FusedGEMMBiasGELU(False, False, M, N, K, 1.0, A, K, B, N, 0.0, Cc, N, bias), 3)
print("fused fp32 FFN GELU: %.3f ms %.1f TF" % (ms, 2.0*M*N*K/ms/1e9))
ms = t(lambda: g.GEMM(False, False, M, N, K, 1.0, A, K, B, N, 0.0, Cc, N), 3)
print("plain fp32 FFN: %.3f ms %.1f TF" % (ms, 2.0*M*N*K/ms/1e9))
PY
