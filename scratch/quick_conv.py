"""Conv2D timings, implicit GEMM vs materialised im2col.  Scratch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0)
e0, e1 = g.Event(), g.Event()
def bench(B, IC, H, W, OC, KH, KW, s, p):
    cp = g.ConvParams(BatchSize=B, InChannels=IC, InHeight=H, InWidth=W, OutChannels=OC, KernelHeight=KH, KernelWidth=KW,
                      StrideH=s, StrideW=s, PadH=p, PadW=p, DilationH=1, DilationW=1, UseBias=True)
    oh, ow = cp.OutputHeight(), cp.OutputWidth()
    n_in, n_w, n_out = B*IC*H*W, OC*IC*KH*KW, B*OC*oh*ow
    x, w, b, o = g.Malloc(n_in*4), g.Malloc(n_w*4), g.Malloc(OC*4), g.Malloc(n_out*4)
    g.FillUniform(x, 1, n_in, n_in, 1); g.FillUniform(w, 1, n_w, n_w, 2); g.FillUniform(b, 1, OC, OC, 3)
    fl = 2.0*OC*IC*KH*KW*B*oh*ow
    res = []
    for r in (0, 1, 2, 3, 4):
        g.SetConvRoute(r)
        for _ in range(2): g.Conv2D(x, w, b, o, cp)
        e0.Record()
        for _ in range(10): g.Conv2D(x, w, b, o, cp)
        e1.Record(); e1.Synchronize()
        ms = g.ElapsedTime(e0, e1)/10
        res.append("%s %.3f ms %.1f TF" % (("auto", "im2col", "t64x128", "t64x64", "t128x128")[r], ms, fl/ms/1e9))
    g.SetConvRoute(-1)
    print("conv B%d %dx%dx%d -> %d k%dx%d s%d p%d: %s" % (B, IC, H, W, OC, KH, KW, s, p, " | ".join(res)), flush=True)
    for q in (x, w, b, o): g.Free(q)
bench(32, 64, 56, 56, 64, 3, 3, 1, 1)
bench(32, 128, 28, 28, 128, 3, 3, 1, 1)
bench(32, 256, 14, 14, 256, 3, 3, 1, 1)
bench(32, 3, 224, 224, 64, 7, 7, 2, 3)
bench(1, 64, 56, 56, 64, 3, 3, 1, 1)
bench(64, 256, 56, 56, 64, 1, 1, 1, 0)
bench(32, 128, 56, 56, 256, 3, 3, 1, 1)
bench(32, 512, 7, 7, 512, 3, 3, 1, 1)
