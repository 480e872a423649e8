"""ncu target / timer for the DMMA DGEMM: usage: prof_dgemm.py N [time]   (time: CUDA-event timing of 3 calls instead of 2 plain calls)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
a, b, c = g.Malloc(n * n * 8), g.Malloc(n * n * 8), g.Malloc(n * n * 8)
L.guda_fill_uniform_f64(a.ptr, n, n, n, 1, 0, 0, n, -1.0, 1.0, None)
L.guda_fill_uniform_f64(b.ptr, n, n, n, 2, 0, 0, n, -1.0, 1.0, None)
def call():
    g.capi.check(L.guda_dgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None), "dgemm")
if len(sys.argv) > 2:
    call(); g.Synchronize()
    e0, e1 = g.Event(), g.Event()
    for rep in range(3):
        e0.Record(); call(); call(); e1.Record(); e1.Synchronize()
        ms = g.ElapsedTime(e0, e1) / 2
        print("dgemm %d^3: %.3f ms  %.2f TFLOP/s" % (n, ms, 2.0 * n ** 3 / ms / 1e9), flush=True)
else:
    call(); call(); g.Synchronize(); print("done", n)
