"""Cooperative-grid kernel, one-barrier form with Phi^-1 in registers / in shared memory (GBDPCG_FLAG_PINV_SMEM) and the two-barrier form\n(GBDPCG_GRID_TWO_BARRIERS=1) on the same systems: kernels, iteration counts, error against the oracle, bit-identity.   usage: python scripts/check_grid_forms.py"""
import os, sys, numpy as np
os.environ.setdefault("GBDPCG_GRID_BPC_SCALE", "1")   # one CTA per SM in both forms: the same decomposition, so the results can be compared bit for bit
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import gbd_pcg_b200 as g
from gbd_pcg_b200 import _lib
from helpers import make_system
from oracle import oracle
import torch
def solve(S,P,gam,lam0,nx,N,tol,mi,flags):
    d=[torch.from_numpy(a).cuda() for a in (S,P,gam)]
    dl=torch.from_numpy(lam0.copy()).cuda(); it=torch.zeros(1,dtype=torch.int32,device="cuda")
    cfg=g.pcg_config(pcg_exit_tol=tol,pcg_max_iter=mi,empty_pinv=0,flags=flags)
    info=g.plan_query(nx,N,S.dtype,True,cfg)
    g.solve_dev_async(nx,N,d[0],d[1],d[2],dl,cfg,it,None); torch.cuda.synchronize()
    return dl.cpu().numpy(), int(it.item()), info["kernel"]
for (nx,nu,N,dt,tol) in [(14,7,1500,np.float32,1e-6),(14,7,1031,np.float32,1e-6),(14,7,1000,np.float64,1e-10),(16,8,1000,np.float64,1e-10),(10,5,2000,np.float64,1e-10)]:
    S,P,gam,nx,N,tol,mi=make_system((nx,nu,N,dt,tol,5000),"stair",seed=6)
    lam0=np.zeros(N*nx,dtype=dt)
    l0,i0,k0=solve(S,P,gam,lam0,nx,N,tol,mi,0)
    l1,i1,k1=solve(S,P,gam,lam0,nx,N,tol,mi,_lib.FLAG_PINV_SMEM)
    os.environ["GBDPCG_GRID_TWO_BARRIERS"]="1"
    l2,i2,k2=solve(S,P,gam,lam0,nx,N,tol,mi,_lib.FLAG_PINV_SMEM)
    del os.environ["GBDPCG_GRID_TWO_BARRIERS"]
    lo,io,_=oracle.pcg(S,P,gam,lam0,nx,N,tol,mi,order=0)
    err=np.linalg.norm(l1.astype(np.float64)-lo)/np.linalg.norm(lo)
    print(nx,N,np.dtype(dt).name,"|",k0,i0,"|",k1,i1,"|",k2,i2,"| oracle",io,"err %.2e"%err,"smem1==reg:",np.array_equal(l0,l1),"smem1==smem2:",np.array_equal(l1,l2))
