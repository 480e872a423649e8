(timeout 900 python -m pytest tests/test_gpu_elementwise.py -m gpu -x -q 2>&1 | tail -12) > gpurun_out/r2a_pytest_ew.log 2>&1
cat gpurun_out/r2a_pytest_ew.log
python - <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import guda_b200 as g
g.Init(0)
e0, e1 = g.Event(), g.Event()
for m, n in ((16384, 16384), (1 << 20, 256), (256, 1 << 20)):
    x, o = g.Malloc(m*n*4), g.Malloc(max(m, n)*4)
    g.FillUniform(x, m, n, n, 1)
    for ax in (0, 1):
        for name, fn in (("sum", g.ReduceSum), ("max", g.ReduceMax)):
            for _ in range(2): fn(x, [m, n], ax, o)
            e0.Record()
            for _ in range(5): fn(x, [m, n], ax, o)
            e1.Record(); e1.Synchronize()
            ms = g.ElapsedTime(e0, e1)/5
            print("reduce_%s %dx%d axis %d: %.3f ms %.0f GB/s" % (name, m, n, ax, ms, 4.0*m*n/ms/1e6), flush=True)
    g.Free(x); g.Free(o)
PY
