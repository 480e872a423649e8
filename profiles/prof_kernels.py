"""One launch of each hot-path kernel at a profiling-friendly size (used under ncu; see profiles/README.md)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g

which = sys.argv[1] if len(sys.argv) > 1 else "all"
g.Init(0)
L = g.load()
reps = int(os.environ.get("PROF_REPS", "3"))

if which in ("all", "level1"):
    n = 1 << 24
    x, y = g.Malloc(n * 4), g.Malloc(n * 4)
    g.FillUniform(x, 1, n, n, 1); g.FillUniform(y, 1, n, n, 2)
    for _ in range(reps):
        g.AXPY(2.5, x, y, n)
        g.DOT(x, y, n)
        g.Scale(0.999, x, n)
    g.Free(x); g.Free(y)
if which in ("all", "level2"):
    m = n = 16384
    a, x, y = g.Malloc(m * n * 4), g.Malloc(n * 4), g.Malloc(n * 4)
    g.FillUniform(a, m, n, n, 11); g.FillUniform(x, 1, n, n, 12); g.FillUniform(y, 1, n, n, 13)
    for _ in range(reps):
        g.Sgemv(False, m, n, 1.0, a, n, x, 1, 0.5, y, 1)
        g.Sgemv(True, m, n, 1.0, a, n, x, 1, 0.5, y, 1)
        g.Sger(m, n, 1e-6, x, 1, y, 1, a, n)
    g.Synchronize()
    for p in (a, x, y):
        g.Free(p)
if which in ("all", "sgemm"):
    n = int(os.environ.get("PROF_N", "8192"))
    a, b, c = g.Malloc(n * n * 4), g.Malloc(n * n * 4), g.Malloc(n * n * 4)
    g.FillUniform(a, n, n, n, 3); g.FillUniform(b, n, n, n, 4)
    for _ in range(reps):
        g.GEMM(False, False, n, n, n, 1.0, a, n, b, n, 0.0, c, n)
    g.Synchronize()
    for p in (a, b, c):
        g.Free(p)
if which in ("all", "fused"):
    M, K, N = 4096, 4096, 8192
    a, b, c, bias = g.Malloc(M * K * 4), g.Malloc(K * N * 4), g.Malloc(M * N * 4), g.Malloc(N * 4)
    g.FillUniform(a, M, K, K, 5); g.FillUniform(b, K, N, N, 6, lo=-0.05, hi=0.05); g.FillUniform(bias, 1, N, N, 7)
    for _ in range(reps):
        g.FusedGEMMBiasGELU(False, False, M, N, K, 1.0, a, K, b, N, 0.0, c, N, bias)
    g.Synchronize()
    for p in (a, b, c, bias):
        g.Free(p)
if which in ("all", "dgemm"):
    n = 4096
    a, b, c = g.Malloc(n * n * 8), g.Malloc(n * n * 8), g.Malloc(n * n * 8)
    g.FillUniform(a, n, n, n, 8, dtype=np.float64); g.FillUniform(b, n, n, n, 9, dtype=np.float64)
    for _ in range(reps):
        g.Dgemm(False, False, n, n, n, 1.0, a, n, b, n, 0.0, c, n)
    g.Synchronize()
    for p in (a, b, c):
        g.Free(p)
if which in ("all", "bf16"):
    M, K, N = 4096, 4096, 8192
    a, b, c, bias = g.Malloc(M * K * 2), g.Malloc(K * N * 2), g.Malloc(M * N * 2), g.Malloc(N * 4)
    g.Memset(a, 0x3c, M * K * 2); g.Memset(b, 0x3c, K * N * 2); g.FillUniform(bias, 1, N, N, 7)
    for _ in range(reps):
        g.FusedGEMMBFloat16BiasGELU(M, N, K, 1.0, a, K, b, N, c, N, bias)
    g.Synchronize()
    for p in (a, b, c, bias):
        g.Free(p)
print("prof_kernels ok:", which)
