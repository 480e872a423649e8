"""ReduceSum / ReduceMax along an axis at 16384 x 16384 (1 GiB), CUDA events over 10 calls."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0)
m = n = 16384
x, o = g.Malloc(m * n * 4), g.Malloc(n * 4)
g.FillUniform(x, m, n, n, 3)
e0, e1 = g.Event(), g.Event()
for name, fn, axis in (("sum axis0", g.ReduceSum, 0), ("sum axis1", g.ReduceSum, 1), ("max axis0", g.ReduceMax, 0), ("max axis1", g.ReduceMax, 1)):
    for _ in range(2): fn(x, [m, n], axis, o)
    e0.Record()
    for _ in range(10): fn(x, [m, n], axis, o)
    e1.Record(); e1.Synchronize()
    ms = g.ElapsedTime(e0, e1) / 10
    print("%s: %.3f ms  %.0f GB/s" % (name, ms, 4.0 * m * n / ms / 1e6))
