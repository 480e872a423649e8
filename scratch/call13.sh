(timeout 900 python -m pytest tests/test_gpu_level2.py -m gpu -x -q -k "dgemv or dger" 2>&1 | tail -12) > gpurun_out/r1k_pytest.log 2>&1
cat gpurun_out/r1k_pytest.log
python - <<'PY'
import ctypes as C, os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
import guda_b200 as g
g.Init(0); L = g.load()
e0, e1 = g.Event(), g.Event()
def timeit(fn, iters=10):
    for i in range(2): fn()
    e0.Record()
    for i in range(iters): fn()
    e1.Record(); e1.Synchronize()
    return g.ElapsedTime(e0, e1) / iters
m = n = 8192
a = g.Malloc(m*n*8); g.FillUniform(a, m, n, n, 7, dtype=np.float64)
x = g.Malloc(n*8); y = g.Malloc(n*8); g.FillUniform(x, 1, n, n, 1, dtype=np.float64); g.FillUniform(y, 1, n, n, 2, dtype=np.float64)
for tr in (0, 1):
    ms = timeit(lambda: L.guda_dgemv(tr, m, n, 1.0, a.ptr, n, x.ptr, 1, 0.5, y.ptr, 1, None))
    print("dgemv %s 8192^2: %.1f us %.0f GB/s" % ("T" if tr else "N", ms*1e3, 8.0*m*n/ms/1e6))
ms = timeit(lambda: L.guda_dger(m, n, 1e-9, x.ptr, 1, y.ptr, 1, a.ptr, n, None))
print("dger 8192^2: %.1f us %.0f GB/s" % (ms*1e3, 16.0*m*n/ms/1e6))
PY
