import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
m = n = int(os.environ.get("MN", "4096"))
a = g.Malloc(m * n * 4); g.FillUniform(a, m, n, n, 7)
x = g.Malloc(n * 4); y = g.Malloc(n * 4); g.FillUniform(x, 1, n, n, 1); g.FillUniform(y, 1, n, n, 2)
for tr in (1, 0):
    for i in range(3):
        L.guda_sgemv(tr, m, n, 1.0, a.ptr, n, x.ptr, 1, 0.5, y.ptr, 1, None)
g.Synchronize()
