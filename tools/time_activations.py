"""GELU / Sigmoid / LinearReLU elementwise timing at n = 2^26 (CUDA events, 20 calls) + the exhaustive division self-test."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
counts, first = (C.c_ulonglong * 2)(), (C.c_uint * 2)()
e0, e1 = g.Event(), g.Event()
e0.Record(); g.capi.check(L.guda_selftest_activations(counts, first), "selftest"); e1.Record(); e1.Synchronize()
print("selftest: gelu mismatches %d (first 0x%08x), sigmoid mismatches %d (first 0x%08x), %.1f ms" % (counts[0], first[0], counts[1], first[1], g.ElapsedTime(e0, e1)))
n = 1 << 26
x = g.Malloc(n * 4); g.FillUniform(x, 1, n, n, 5, lo=-4.0, hi=4.0)
for name, fn in (("gelu", lambda: g.GELU(x, n)), ("sigmoid", lambda: g.Sigmoid(x, n)), ("linear_relu", lambda: g.LinearReLU(x, 1.01, 0.1, n))):
    g.FillUniform(x, 1, n, n, 5, lo=-4.0, hi=4.0)
    for _ in range(3): fn()
    e0.Record()
    for _ in range(20): fn()
    e1.Record(); e1.Synchronize()
    ms = g.ElapsedTime(e0, e1) / 20
    print("%s: %.3f ms  %.0f GB/s" % (name, ms, 8.0 * n / ms / 1e6))
