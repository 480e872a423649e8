"""Quick SGEMM timings (exact route) + FFMA/FFMA2 peaks + cuBLAS fp32 calibration.  Scratch, not a bench."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
def peak(k):
    v = C.c_double(0); L.guda_microbench_peak(k, 3, C.byref(v)); return v.value
print("FFMA %.1f  FFMA2 %.1f  DFMA %.1f  DMMA %.1f TFLOP/s" % (peak(0), peak(4), peak(1), peak(3)))
e0, e1 = g.Event(), g.Event()
for n, iters in ((1024, 50), (2048, 20), (4096, 5), (8192, 3)):
    a, b, c = g.Malloc(n*n*4), g.Malloc(n*n*4), g.Malloc(n*n*4)
    g.FillUniform(a, n, n, n, 1); g.FillUniform(b, n, n, n, 2)
    for _ in range(3): L.guda_sgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None)
    e0.Record()
    for _ in range(iters): L.guda_sgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None)
    e1.Record(); e1.Synchronize()
    ms = g.ElapsedTime(e0, e1) / iters
    print("sgemm %d^3: %.3f ms  %.1f TFLOP/s" % (n, ms, 2.0*n**3/ms/1e9))
    for p in (a, b, c): g.Free(p)
if len(sys.argv) > 1 and sys.argv[1] == "cublas":
    import torch
    torch.backends.cuda.matmul.allow_tf32 = False
    for n in (1024, 8192):
        x = torch.randn(n, n, device="cuda"); y = torch.randn(n, n, device="cuda")
        for _ in range(3): x @ y
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(5): x @ y
        e.record(); torch.cuda.synchronize()
        print("cuBLAS fp32 (no tf32) %d^3: %.1f TFLOP/s" % (n, 2.0*n**3*5/s.elapsed_time(e)/1e9))
