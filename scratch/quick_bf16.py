"""bf16 tensor-core GEMM timings at the C3 FFN shape (plain + fused GELU) and 8192^3.  Scratch."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import guda_b200 as g
g.Init(0); L = g.load()
e0, e1 = g.Event(), g.Event()
def t(fn, iters=20):
    for _ in range(3): fn()
    e0.Record()
    for _ in range(iters): fn()
    e1.Record(); e1.Synchronize()
    return g.ElapsedTime(e0, e1) / iters
for (M, K, N) in ((8192, 4096, 16384), (8192, 8192, 8192), (4096, 4096, 4096)):
    a, b, c, bias = g.Malloc(M*K*2), g.Malloc(K*N*2), g.Malloc(M*N*2), g.Malloc(N*4)
    tmp = g.Malloc(max(M*K, K*N)*4)
    s = 1.0/np.sqrt(K)
    g.FillUniform(tmp, M, K, K, 510); L.guda_f32_to_bf16(tmp.ptr, a.ptr, M*K, None)
    g.FillUniform(tmp, K, N, N, 511, lo=-3*s, hi=3*s); L.guda_f32_to_bf16(tmp.ptr, b.ptr, K*N, None)
    g.FillUniform(bias, 1, N, N, 512); g.Synchronize(); g.Free(tmp)
    ms = t(lambda: L.guda_gemm_bf16(0, 0, M, N, K, 1.0, a.ptr, K, b.ptr, N, 0.0, c.ptr, N, None))
    ms2 = t(lambda: L.guda_gemm_bf16_bias_gelu(M, N, K, 1.0, a.ptr, K, b.ptr, N, c.ptr, N, bias.ptr, None))
    print("bf16 %dx%dx%d: plain %.3f ms %.0f TF/s   fused-gelu %.3f ms %.0f TF/s" % (M, K, N, ms, 2.0*M*N*K/ms/1e9, ms2, 2.0*M*N*K/ms2/1e9))
    for p in (a, b, c, bias): g.Free(p)
