"""ncu target for the headline kernel: ONE SGEMM n^3 call (default 65536: 8 K-chunks = 8 pre-pass + 8 pipeline launches)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
a, b, c = g.Malloc(n * n * 4), g.Malloc(n * n * 4), g.Malloc(n * n * 4)
g.FillUniform(a, n, n, n, 11); g.FillUniform(b, n, n, n, 12)
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 1):
    g.capi.check(L.guda_sgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None), "sgemm")
g.Synchronize()
print("done", n)
