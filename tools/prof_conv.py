import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import guda_b200 as g
g.Init(0)
B, IC, H, W, OC = 32, 64, 56, 56, 64
cp = g.ConvParams(BatchSize=B, InChannels=IC, InHeight=H, InWidth=W, OutChannels=OC, KernelHeight=3, KernelWidth=3,
                  StrideH=1, StrideW=1, PadH=1, PadW=1, DilationH=1, DilationW=1, UseBias=True)
n_in, n_w, n_out = B*IC*H*W, OC*IC*9, B*OC*H*W
x, w, b, o = g.Malloc(n_in*4), g.Malloc(n_w*4), g.Malloc(OC*4), g.Malloc(n_out*4)
g.FillUniform(x, 1, n_in, n_in, 1); g.FillUniform(w, 1, n_w, n_w, 2); g.FillUniform(b, 1, OC, OC, 3)
for _ in range(3): g.Conv2D(x, w, b, o, cp)
