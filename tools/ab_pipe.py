"""A/B harness for the exact SGEMM kernels: (1) LDS wavefront cost per lane pattern, (2) bit-equality of every sgemm_pipe configuration against the
register-staged kernel, (3) timing table per configuration and size.  Scratch (not part of the product or the tests)."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0); L = g.load()
e0, e1 = g.Event(), g.Event()
what = sys.argv[1] if len(sys.argv) > 1 else "all"

CFGS = [int(v) for v in os.environ.get("PIPE_CFGS", "0,1,2,3").split(",")]

if what in ("all", "check"):
    # bit-equality against the register-staged kernel (cfg 0), incl. alpha != 1, beta != 0, zeros in A, TN form
    rng = np.random.default_rng(5)
    for (m, n, k, ta, alpha, beta, zeros) in [(256, 256, 256, 0, 1.0, 0.0, False), (512, 256, 320, 0, 1.5, 0.5, False), (256, 512, 128, 1, 0.75, 1.0, False),
                                             (512, 512, 256, 0, 1.0, 0.0, True), (256, 256, 48, 0, 2.0, 0.25, True), (1024, 1024, 1024, 0, 1.0, 0.0, False)]:
        a = (rng.random((k, m) if ta else (m, k), dtype=np.float32) * 2 - 1)
        b = (rng.random((k, n), dtype=np.float32) * 2 - 1)
        c0 = (rng.random((m, n), dtype=np.float32) * 2 - 1)
        if zeros:
            a[rng.random(a.shape) < 0.01] = 0.0
            b[3, 5] = np.inf; b[7, 9] = np.nan
            c0[c0 < -0.9] = -0.0
        da, db = g.to_device(a), g.to_device(b)
        ref = None
        for cfg in CFGS:
            g.capi.check(L.guda_set_sgemm_pipe(cfg, -1), "pipe")
            dc = g.to_device(c0)
            g.capi.check(L.guda_sgemm(ta, 0, m, n, k, alpha, da.ptr, m if ta else k, db.ptr, n, beta, dc.ptr, n, None), "sgemm")
            out = dc.Float32().reshape(m, n).copy()
            g.Free(dc)
            if cfg == 0:
                ref = out
            else:
                same = np.array_equal(out.view(np.uint32), ref.view(np.uint32))
                print("check m=%d n=%d k=%d ta=%d alpha=%g beta=%g zeros=%d cfg=%d: %s" % (m, n, k, ta, alpha, beta, zeros, cfg,
                      "bit-identical" if same else "MISMATCH (%d elements)" % int((out.view(np.uint32) != ref.view(np.uint32)).sum())), flush=True)
        g.Free(da); g.Free(db)

if what in ("all", "time"):
    sizes = [int(v) for v in sys.argv[2:]] or [1024, 2048, 4096, 8192]
    for n in sizes:
        nset = 4 if n <= 4096 else 2
        sets = []
        for i in range(nset):
            a, b, c = g.Malloc(n * n * 4), g.Malloc(n * n * 4), g.Malloc(n * n * 4)
            g.FillUniform(a, n, n, n, 1 + i); g.FillUniform(b, n, n, n, 20 + i)
            sets.append((a, b, c))
        iters = max(3, min(100, int(1.2e11 / (2.0 * n ** 3))))

        def run(i):
            a, b, c = sets[i % nset]
            g.capi.check(L.guda_sgemm(0, 0, n, n, n, 1.0, a.ptr, n, b.ptr, n, 0.0, c.ptr, n, None), "sgemm")
        for cfg in CFGS:
            for lag in ((1,) if cfg == 0 else [int(v) for v in os.environ.get("PIPE_LAGS", "1").split(",")]):
                g.capi.check(L.guda_set_sgemm_pipe(cfg, lag), "pipe")
                for i in range(3):
                    run(i)
                best = 1e9
                for rep in range(3):
                    e0.Record()
                    for i in range(iters):
                        run(i)
                    e1.Record(); e1.Synchronize()
                    best = min(best, g.ElapsedTime(e0, e1) / iters)
                print("time cfg=%d lag=%d sgemm %d^3: %.2f us  %.2f TFLOP/s" % (cfg, lag, n, best * 1e3, 2.0 * n ** 3 / best / 1e9), flush=True)
        for s in sets:
            for p in s:
                g.Free(p)
    g.capi.check(L.guda_set_sgemm_pipe(-1, 1), "pipe")
