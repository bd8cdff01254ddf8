"""Host-side replay of the multi-matrix producer (mia_tile_multi.cuh): for the 13 probes of an optimiser iteration, how many
matrices share the first box of a tile position, how many ring stages a position needs with up to K boxes, and how many
matrices are left for the per-voxel gather.  CPU only.  Usage: python scripts/sim_multi_boxes.py"""
import numpy as np, sys
sys.path.insert(0,'/root/repo')
from medicalimageanalysis_b200 import geometry as G
from medicalimageanalysis_b200.registration import probe_params
def sim(dims, sp, BOX, xalign, base=(0,0,0,0,0,0), h=(0.5,0.5), K=3):
    nx,ny,nz = dims
    g = G.Grid(dims, sp, tuple(-0.5*np.array(sp)*np.array(dims)))
    center = np.zeros(3)
    p2w = G.index_to_physical(g); w2p = G.physical_to_index(g)
    deltas = np.array([h[0]]*3+[h[1]]*3)
    As = [ (w2p @ G.params_to_matrix(p, center) @ p2w)[:3] for p in probe_params(np.array(base,float), 1.0, deltas)]
    TX,TY,TZ = 32,8,8
    ext = np.array([TX-1,TY-1,TZ-1.0])
    # sample tiles on a coarse lattice
    tix = np.arange(0, nx, TX); tiy = np.arange(0, ny, TY); tiz = np.arange(0, nz, TZ)
    I,J,Kk = np.meshgrid(tix[::2], tiy[::4], tiz[::4], indexing='ij')
    o = np.stack([I.ravel(),J.ravel(),Kk.ravel()],1).astype(float)
    lo=[];hi=[]
    n=np.array([nx,ny,nz])
    for A in As:
        c0 = o @ A[:,:3].T + A[:,3]
        d = A[:,:3]*ext[None,:]
        neg = np.minimum(d,0).sum(1); pos = np.maximum(d,0).sum(1)
        l = np.clip(c0+neg, 0, n-1)-1e-3; hh = np.clip(c0+pos,0,n-1)+1e-3
        lo.append(np.floor(l).astype(int)); hi.append(np.floor(hh).astype(int)+2)
    lo=np.stack(lo,1); hi=np.stack(hi,1)   # [tiles, 13, 3]
    T=lo.shape[0]
    B=np.array(BOX)
    stages=np.zeros(T,int); gather=np.zeros(T,int); fit1=np.zeros(T,int)
    for t in range(T):
        rem=list(range(13)); k=0
        while rem and k<K:
            # union origin
            u=lo[t,rem].min(0).copy(); u[0]&=~(xalign-1)
            fu=[m for m in rem if (lo[t,m]>=u).all() and (hi[t,m]-u<=B).all()]
            a=lo[t,rem[0]].copy(); a[0]&=~(xalign-1)
            fa=[m for m in rem if (lo[t,m]>=a).all() and (hi[t,m]-a<=B).all()]
            f = fu if len(fu)>=len(fa) else fa
            if not f: break
            if k==0: fit1[t]=len(f)
            rem=[m for m in rem if m not in f]; k+=1
        stages[t]=max(k,1); gather[t]=len(rem)
    print(dims, BOX, "K=%d"%K, "mean fit in box1 %.2f/13"%fit1.mean(), "mean stages %.2f"%stages.mean(), "mean gathered matrices %.2f"%gather.mean())
for K in (1,2,3):
    sim((1024,1024,1200),(0.5,0.5,0.5),(96,15,12),8,K=K)
sim((1024,1024,1200),(0.5,0.5,0.5),(96,15,14),8,K=1)
sim((1024,1024,1200),(0.5,0.5,0.5),(96,15,14),8,K=2)
sim((512,512,100),(0.9765625,0.9765625,2.5),(96,15,12),8,base=(3,-2,4,5,-3,7),K=1)
sim((256,256,256),(1,1,1),(64,15,14),4,base=(4,3,-5,6,-4,3),K=1)
sim((256,256,256),(1,1,1),(64,15,14),4,base=(4,3,-5,6,-4,3),K=2)
