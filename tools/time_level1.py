#!/usr/bin/env python
"""Level-1 / reduction timing loop (CUDA events on the library's stream, rotating buffers larger than L2), several repeats so the
box-to-box and run-to-run spread is visible:  python tools/time_level1.py [n_log2=24] [repeats=5]
Prints GB/s for SAXPY, SDOT (device result, back to back), the no-allocation SDOT, Max, and Min of a post-ReLU vector (the
streaming pass must not fall back to the single-CTA ordered replay when only +0 occurs)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g  # noqa: E402


def main():
    lg = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    n = 1 << lg
    g.Init(0)
    L = g.load()
    npairs = max(4, (512 << 20) // (8 * n))
    pairs = []
    for i in range(npairs):
        x, y = g.Malloc(n * 4), g.Malloc(n * 4)
        g.FillUniform(x, 1, n, n, 100 + i); g.FillUniform(y, 1, n, n, 200 + i)
        pairs.append((x, y))
    relu = g.Malloc(n * 4)
    g.FillUniform(relu, 1, n, n, 77)
    g.ReLU(relu, n) if hasattr(g, "ReLU") else None
    scal = g.Malloc(64)
    e0, e1 = g.Event(), g.Event()
    it = [0]

    def timed(fn, iters):
        for _ in range(8):
            fn()
        e0.Record()
        for _ in range(iters):
            fn()
        e1.Record(); e1.Synchronize()
        return g.ElapsedTime(e0, e1) / iters

    def axpy():
        x, y = pairs[it[0] % npairs]; it[0] += 1
        g.AXPY(2.5, x, y, n)

    def dot():
        x, y = pairs[it[0] % npairs]; it[0] += 1
        L.guda_sdot_device(n, x.ptr, 1, y.ptr, 1, scal.ptr, None)

    def dot_static():
        x, y = pairs[it[0] % npairs]; it[0] += 1
        L.guda_sdot_device_static(n, x.ptr, 1, y.ptr, 1, scal.ptr, None)

    def vmax():
        x, _ = pairs[it[0] % npairs]; it[0] += 1
        x.Max(n)

    def vmin_relu():
        relu.Min(n)

    for r in range(reps):
        out = {}
        out["saxpy"] = 12.0 * n / (timed(axpy, 40) * 1e-3) / 1e9
        out["sdot_device"] = 8.0 * n / (timed(dot, 40) * 1e-3) / 1e9
        out["sdot_static"] = 8.0 * n / (timed(dot_static, 40) * 1e-3) / 1e9
        out["smax_blocking"] = 4.0 * n / (timed(vmax, 20) * 1e-3) / 1e9
        out["smin_post_relu_blocking"] = 4.0 * n / (timed(vmin_relu, 20) * 1e-3) / 1e9
        print("n=2^%d rep %d: " % (lg, r) + "  ".join("%s %.0f GB/s" % kv for kv in out.items()), flush=True)


if __name__ == "__main__":
    main()
