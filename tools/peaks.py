import sys, ctypes as C
sys.path.insert(0,'/root/repo')
import guda_b200 as g
g.Init(0); L=g.load()
for kind,name in [(0,'FFMA TF'),(1,'DFMA TF'),(3,'DMMA TF'),(2,'copy GB/s')]:
    v=C.c_double(0); rc=L.guda_microbench_peak(kind,3,C.byref(v)); print(name, rc, round(v.value,2))
