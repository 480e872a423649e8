"""Quick Level-1 timings over rotating buffers (512 MiB > L2).  Scratch."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
n = 1 << 24
pairs = []
for i in range(4):
    x, y = g.Malloc(n*4), g.Malloc(n*4)
    g.FillUniform(x, 1, n, n, 100+i); g.FillUniform(y, 1, n, n, 200+i); pairs.append((x, y))
scal = g.Malloc(64)
e0, e1 = g.Event(), g.Event()
def timeit(fn, iters=200):
    for i in range(8): fn(i)
    e0.Record()
    for i in range(iters): fn(i)
    e1.Record(); e1.Synchronize()
    return g.ElapsedTime(e0, e1) / iters
ms = timeit(lambda i: L.guda_sdot_device(n, pairs[i%4][0].ptr, 1, pairs[i%4][1].ptr, 1, scal.ptr, None))
print("sdot  16M: %.2f us  %.0f GB/s" % (ms*1e3, 8.0*n/ms/1e6))
ms = timeit(lambda i: L.guda_saxpy(n, 2.5, pairs[i%4][0].ptr, 1, pairs[i%4][1].ptr, 1, None))
print("saxpy 16M: %.2f us  %.0f GB/s" % (ms*1e3, 12.0*n/ms/1e6))
ms = timeit(lambda i: L.guda_sscal(n, 0.999, pairs[i%4][0].ptr, 1, None))
print("sscal 16M: %.2f us  %.0f GB/s" % (ms*1e3, 8.0*n/ms/1e6))
