"""One launch of each kernel added late in round 1, at the bench sizes (for ncu --set full)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0)
ne = 1 << 26
a, b, c = g.Malloc(ne*4), g.Malloc(ne*4), g.Malloc(ne*4)
g.FillUniform(a, 1, ne, ne, 1); g.FillUniform(b, 1, ne, ne, 2)
for _ in range(2):
    g.Add(a, b, c, ne); g.LinearReLU(c, 1.01, 0.1, ne); g.GELU(c, ne); g.Sigmoid(c, ne)
for p in (a, b, c): g.Free(p)
m = n = 16384
x, o = g.Malloc(m*n*4), g.Malloc(n*4)
g.FillUniform(x, m, n, n, 3)
for _ in range(2):
    g.ReduceSum(x, [m, n], 1, o); g.ReduceMax(x, [m, n], 0, o); g.ReduceSum(x, [m, n], 0, o)
g.Free(x); g.Free(o)
B, IC, H, W, OC = 32, 64, 56, 56, 64
cp = g.ConvParams(BatchSize=B, InChannels=IC, InHeight=H, InWidth=W, OutChannels=OC, KernelHeight=3, KernelWidth=3,
                  StrideH=1, StrideW=1, PadH=1, PadW=1, DilationH=1, DilationW=1, UseBias=True)
n_in, n_w, n_out = B*IC*H*W, OC*IC*9, B*OC*H*W
x, w, bb, o = g.Malloc(n_in*4), g.Malloc(n_w*4), g.Malloc(OC*4), g.Malloc(n_out*4)
g.FillUniform(x, 1, n_in, n_in, 1); g.FillUniform(w, 1, n_w, n_w, 2); g.FillUniform(bb, 1, OC, OC, 3)
for _ in range(2): g.Conv2D(x, w, bb, o, cp)
cp2 = g.ConvParams(BatchSize=32, InChannels=128, InHeight=56, InWidth=56, OutChannels=256, KernelHeight=3, KernelWidth=3,
                   StrideH=1, StrideW=1, PadH=1, PadW=1, DilationH=1, DilationW=1, UseBias=True)
x2, w2, b2, o2 = g.Malloc(32*128*56*56*4), g.Malloc(256*128*9*4), g.Malloc(256*4), g.Malloc(32*256*56*56*4)
g.FillUniform(x2, 1, 32*128*56*56, 32*128*56*56, 4); g.FillUniform(w2, 1, 256*128*9, 256*128*9, 5); g.FillUniform(b2, 1, 256, 256, 6)
for _ in range(2): g.Conv2D(x2, w2, b2, o2, cp2)
n = 1024
A, Bm, C_ = g.Malloc(n*n*4), g.Malloc(n*n*4), g.Malloc(n*n*4)
g.FillUniform(A, n, n, n, 7); g.FillUniform(Bm, n, n, n, 8)
for _ in range(2): g.GEMM(False, False, n, n, n, 1.0, A, n, Bm, n, 0.0, C_, n)
g.Synchronize()
