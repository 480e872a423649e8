import numpy as np, sys
sys.path.insert(0,'/root/repo')
import guda_b200 as g
g.Init(0)
for (m,n,k) in [(128,128,32),(128,128,64),(128,256,32)]:
    rng=np.random.default_rng(0)
    a=rng.integers(-4,5,size=(m,k)).astype(np.float32); b=rng.integers(-4,5,size=(k,n)).astype(np.float32)
    c0=np.zeros((m,n),np.float32)
    da,db,dc=g.to_device(a.ravel()),g.to_device(b.ravel()),g.to_device(c0.ravel())
    g.SetSgemmMode(1)
    g.GEMM(False,False,m,n,k,1.0,da,k,db,n,0.0,dc,n); g.Synchronize()
    g.SetSgemmMode(0)
    got=dc.Float32().reshape(m,n); want=a@b
    print((m,n,k),'max err',np.abs(got-want).max(),'nonzero got',np.count_nonzero(got),'of',got.size)
    print(' got row0',got[0,:8]); print(' want row0',want[0,:8])
    # which partial sums match? try k-subsets
    for kk in (8,16,24,32):
        if kk<=k: print('  match first',kk,'k:',np.allclose(got,a[:,:kk]@b[:kk,:]))
    # column permutation check
    for sh in (32,64):
        if np.allclose(got[:, :32], want[:, sh:sh+32]): print('  cols shifted by',sh)
