"""C1 dissection: 30 back-to-back SGEMM 1024^3 calls (for an ncu launch list) and the same timed with events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
n = 1024
sets = []
for i in range(8):
    a, b, c = g.Malloc(n * n * 4), g.Malloc(n * n * 4), g.Malloc(n * n * 4)
    g.FillUniform(a, n, n, n, 1 + i); g.FillUniform(b, n, n, n, 20 + i)
    sets.append((a, b, c))
e0, e1 = g.Event(), g.Event()
import time
for it in range(3):
    g.Synchronize()
    t0 = time.perf_counter()
    e0.Record()
    for i in range(30):
        a, b, c = sets[i % 8]
        L.guda_sgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None)
    t1 = time.perf_counter()
    e1.Record(); e1.Synchronize()
    print("sgemm 1024^3: %.2f us per call on the device, %.2f us per call of host time to enqueue" % (g.ElapsedTime(e0, e1) / 30 * 1e3, (t1 - t0) / 30 * 1e6))
