"""Quick Level-2 timings (matrix >> L2).  Scratch."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
e0, e1 = g.Event(), g.Event()
def timeit(fn, iters=20):
    for i in range(3): fn(i)
    e0.Record()
    for i in range(iters): fn(i)
    e1.Record(); e1.Synchronize()
    return g.ElapsedTime(e0, e1) / iters
for m, n in [(16384, 16384), (4096, 4096), (8192, 8192), (8192, 32768), (65536, 4096), (4096, 65536), (16384, 2048), (16384, 1024)]:
    nbuf = max(2, min(4, (1 << 30) // (m * n * 4) + 1))
    mats = []
    for i in range(nbuf):
        a = g.Malloc(m * n * 4); g.FillUniform(a, m, n, n, 7 + i); mats.append(a)
    x = g.Malloc(max(m, n) * 4); y = g.Malloc(max(m, n) * 4)
    g.FillUniform(x, 1, max(m, n), max(m, n), 1); g.FillUniform(y, 1, max(m, n), max(m, n), 2)
    for rows in ("",):
        if rows: os.environ["GUDA_GEMV_ROWS"] = rows
        else: os.environ.pop("GUDA_GEMV_ROWS", None)
        ms = timeit(lambda i: L.guda_sgemv(0, m, n, 1.0, mats[i % nbuf].ptr, n, x.ptr, 1, 0.5, y.ptr, 1, None))
        print("sgemv N %6dx%-6d rows/warp=%-4s %.1f us  %.0f GB/s" % (m, n, rows or "auto", ms * 1e3, 4.0 * m * n / ms / 1e6), flush=True)
    for cfg in ("1", "2", "0"):
        os.environ["GUDA_GEMVT_COLS"] = cfg
        ms = timeit(lambda i: L.guda_sgemv(1, m, n, 1.0, mats[i % nbuf].ptr, n, x.ptr, 1, 0.5, y.ptr, 1, None))
        print("sgemv T %6dx%-6d cols/lane=%s %.1f us  %.0f GB/s" % (m, n, cfg, ms * 1e3, 4.0 * m * n / ms / 1e6), flush=True)
    os.environ.pop("GUDA_GEMVT_COLS")
    ms = timeit(lambda i: L.guda_sger(m, n, 1e-6, x.ptr, 1, y.ptr, 1, mats[i % nbuf].ptr, n, None))
    print("sger    %6dx%-6d %.1f us  %.0f GB/s" % (m, n, ms * 1e3, 8.0 * m * n / ms / 1e6), flush=True)
    for a in mats: g.Free(a)
    g.Free(x); g.Free(y)
