import numpy as np, torch, sys, time, os, ctypes as C
sys.path.insert(0,".")
import sr459
from sr459 import _lib
from sr459.bands import BandPlan, run_emulated_peer
from oracle import sr_oracle as O
NOM=[(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
rng=np.random.default_rng(32); h,w=150,168
lrs=[rng.uniform(0,255,(h,w)).astype(np.float32).astype(np.float64) for _ in NOM]
psf=O.make_gaussian_psf()
class V:
    def __init__(s,p,n): s.__cuda_array_interface__={"shape":(n,),"typestr":"<u4","data":(int(p),False),"version":2}
def flags(p):
    ptr=C.c_void_p(); _lib.check(p._lib.sr459_band_export(p._h,None,C.byref(ptr)))
    return torch.as_tensor(V(ptr.value,8),device="cuda").cpu().tolist()
world=int(sys.argv[1]) if len(sys.argv)>1 else 2
plans=[BandPlan(4,h,w,NOM,psf,r,world) for r in range(world)]
for K in (0,4):
    t=time.time()
    try:
        hr,e=run_emulated_peer(plans,lrs,K); print(world,K,"ok",round(time.time()-t,3))
    except Exception as ex:
        print(world,K,"fail",round(time.time()-t,3),str(ex)[:60])
    torch.cuda.synchronize()
    for p in plans: print("  flags [fromUp,fromDown,err,base,ticket,cntUp,cntDown]", flags(p))
