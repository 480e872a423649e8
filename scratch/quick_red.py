"""Quick reduction timings (512 MiB rotating > L2).  Scratch."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
n = 1 << 24
bufs = []
for i in range(8):
    x = g.Malloc(n*4); g.FillUniform(x, 1, n, n, 100+i); bufs.append(x)
rf, ri = C.c_float(0), C.c_longlong(0)
for name, fn, res in (("ssum", L.guda_ssum, rf), ("ssumsq", L.guda_ssumsq, rf), ("smax", L.guda_smax, rf), ("smin", L.guda_smin, rf),
                      ("sargmax", L.guda_sargmax, ri), ("sargmin", L.guda_sargmin, ri)):
    for i in range(8): fn(n, bufs[i % 8].ptr, C.byref(res), None)
    t0 = time.perf_counter()
    for i in range(200): fn(n, bufs[i % 8].ptr, C.byref(res), None)
    dt = (time.perf_counter() - t0) / 200
    print("%-8s 16M blocking call: %.1f us  %.0f GB/s" % (name, dt*1e6, 4.0*n/dt/1e9), flush=True)
for i in range(8): L.guda_isamax(n, bufs[i % 8].ptr, 1, C.byref(ri), None)
t0 = time.perf_counter()
for i in range(200): L.guda_isamax(n, bufs[i % 8].ptr, 1, C.byref(ri), None)
dt = (time.perf_counter() - t0) / 200
print("isamax   16M blocking call: %.1f us  %.0f GB/s" % (dt*1e6, 4.0*n/dt/1e9))
