import sys, numpy as np
sys.path.insert(0,'.')
import oracle
from stencil_b200 import capi, sweep
capi.check(capi.lib().sb200_init(0))
for dt,code in ((np.float64,oracle.F64),(np.float32,oracle.F32)):
  for n0 in (4,5,6,7,8,9):
    for n1 in (4,5):
        n=(n0,n1,3); T=2; k=1
        a=oracle.fill(n,1,code,oracle.INIT_RANDOM,1); b=oracle.fill(n,1,code,oracle.INIT_REF)
        ga,gb=a.copy(),b.copy()
        want = b if oracle.sweep(n,1,oracle.BOX,T,a,b) else a
        got,t=sweep.run_host(n,ga,gb,T,radius=1,shape=capi.BOX,time_block=k,kernel=capi.KERNEL_TMA3D)
        d=np.abs(want.astype(np.float64)-got.astype(np.float64))
        idx=np.argwhere(d>0)
        print(np.dtype(dt).name,n,"nbad",(d>0).sum(),"x:",sorted(set(idx[:,2].tolist())),"y:",sorted(set(idx[:,1].tolist())))
