"""SGEMM exact route, small/medium sizes, per forced tile (GUDA_SIMT_TILE).  Scratch."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
e0, e1 = g.Event(), g.Event()
sizes = [int(v) for v in sys.argv[1:]] or [512, 1024, 1536, 2048, 3072, 4096]
for n in sizes:
    sets = []
    for i in range(4):
        a, b, c = g.Malloc(n*n*4), g.Malloc(n*n*4), g.Malloc(n*n*4)
        g.FillUniform(a, n, n, n, 1 + i); g.FillUniform(b, n, n, n, 20 + i)
        sets.append((a, b, c))
    iters = max(4, min(200, int(2e11 / (2.0 * n**3))))
    def run(i):
        a, b, c = sets[i & 3]
        L.guda_sgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None)
    for i in range(4): run(i)
    best = 1e9
    for rep in range(3):
        e0.Record()
        for i in range(iters): run(i)
        e1.Record(); e1.Synchronize()
        best = min(best, g.ElapsedTime(e0, e1) / iters)
    print("tile=%s sgemm %d^3: %.2f us  %.1f TFLOP/s" % (os.environ.get("GUDA_SIMT_TILE", "auto"), n, best*1e3, 2.0*n**3/best/1e9), flush=True)
    for s in sets:
        for p in s: g.Free(p)
