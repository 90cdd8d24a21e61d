import numpy as np, sys
sys.path.insert(0,'.')
import tci_b200 as tci
from oracle import oracle as O
h=tci.default_handle(); dev="cuda:0"
rng=np.random.default_rng(1)
for q,r in [(30,7),(40,40),(100,33)]:
    P=rng.standard_normal((r,r))+3*np.eye(r); B=rng.standard_normal((q,r))
    X=tci.solve_right(h,tci.Mat.from_numpy(B,dev),tci.Mat.from_numpy(P,dev)).numpy()
    ref=B@np.linalg.inv(P)
    d=np.abs(X-ref)
    print(q,r,'maxdiff',d.max(),'bad cols',np.where(d.max(axis=0)>1e-9)[0][:20])
# trsm direct
from tci_b200._lib import lib
import ctypes as C
for r,q,lower,unit in [(7,5,1,1),(7,5,0,0),(40,10,1,1),(40,10,0,0),(70,9,1,0)]:
    T=rng.standard_normal((r,r))+4*np.eye(r)
    T=np.tril(T) if lower else np.triu(T)
    Tu=T.copy()
    if unit: np.fill_diagonal(Tu,1.0)
    B=rng.standard_normal((q,r))
    Tm=tci.Mat.from_numpy(T,dev); Bm=tci.Mat.from_numpy(B,dev)
    h.check(lib().tci_trsm_right_f64(h.h,Tm.ptr,C.c_int64(1),C.c_int64(Tm.ld),C.c_int64(r),lower,unit,Bm.ptr,C.c_int64(1),C.c_int64(Bm.ld),C.c_int64(q)))
    X=Bm.numpy(); ref=B@np.linalg.inv(Tu)
    d=np.abs(X-ref); print('trsm',r,q,lower,unit,d.max(),np.where(d.max(axis=0)>1e-9)[0][:20])
f3 = tci.SumFunction([3, 3, 3, 3])
try:
    print(f3([[1], [2]], [[1], [2]], 1).shape)
    print(f3([], [[1]], 2).shape)
    f3([[1]], [[1]], 1)
except Exception as e: print('EXC',repr(e))
