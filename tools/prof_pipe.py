"""ncu target: a few launches of one sgemm_pipe configuration.  usage: r2_prof_pipe.py CFG N [LAG]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
cfg, n = int(sys.argv[1]), int(sys.argv[2])
lag = int(sys.argv[3]) if len(sys.argv) > 3 else 1
a, b, c = g.Malloc(n * n * 4), g.Malloc(n * n * 4), g.Malloc(n * n * 4)
g.FillUniform(a, n, n, n, 1); g.FillUniform(b, n, n, n, 2)
g.capi.check(L.guda_set_sgemm_pipe(cfg, lag), "pipe")
for _ in range(3):
    g.capi.check(L.guda_sgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None), "sgemm")
g.Synchronize()
print("done cfg", cfg, "n", n)
