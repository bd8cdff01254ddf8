"""Shared-memory bank simulator behind the box pitches of csrc/mia_tile.cuh and csrc/mia_tile_x2.cuh (CPU only).

For a warp whose 32 lanes sample x-neighbours (xstride 1: lane = voxel) or x-pairs (xstride 2: lane = two voxels) of an
output row through a rigid matrix, count the wavefronts of each of the 8 tap loads: the largest number of distinct
32-bit words any bank has to serve.  PX / PY = box row length (elements) and rows per plane, es = element size.
ncu on B200 agrees with it: 1.27 wavefronts per LDS for 96 / xstride 1, 1.015 for 128 / xstride 2 on the config-2 matrix."""
import numpy as np
A=np.array([[0.996956,-0.071483,-0.079657,3.97465],[0.069714,0.99607,-0.139877,-33.989834],[0.013633,0.020431,0.998021,-17.830058]])
def wavefronts(word):
    bank=word%32; mx=0
    for b in np.unique(bank):
        mx=max(mx,len(np.unique(word[bank==b])))
    return mx
def sim(PX,PY,xstride,Amat,ntile=150,seed=0,es=2):
    rng=np.random.default_rng(seed); tot=0;n=0
    for t in range(ntile):
        i0=rng.integers(0,8)*64; j0=rng.integers(8,60)*8; k0=rng.integers(4,30)*8
        for w in range(8):
            for k in range(0,8,3):
              for sub in range(xstride):
                i=i0+np.arange(32)*xstride+sub; j=j0+w; kk=k0+k
                c=Amat[:,0:1]*i+Amat[:,1:2]*j+Amat[:,2:3]*kk+Amat[:,3:4]
                f=np.floor(c).astype(int)
                for dx in (0,1):
                    for dy in (0,1):
                        for dz in (0,1):
                            e=(f[0]+dx)+(f[1]+dy)*PX+(f[2]+dz)*PX*PY
                            tot+=wavefronts((e*es)//4);n+=1
    return tot/n
def rot(rx,ry,rz):
    from scipy.spatial.transform import Rotation as R
    M=R.from_euler('xyz',[rx,ry,rz],degrees=True).as_matrix()
    return np.hstack([M,np.array([[3.3],[40.2],[20.1]])])
print("C2 matrix")
for PX,xs in ((96,1),(64,1),(128,2),(96,2),(80,2),(64,2),(72,2),(88,2),(112,2)):
    print(PX,"xstride",xs, [round(sim(PX,PY,xs,A),3) for PY in (14,15,16)])
for ang in ((0.3,0.2,-0.4),(5,5,5),(0,0,10),(8,-6,3),(-3,2,-4)):
    Am=rot(*ang)
    print(ang,"96/1:",round(sim(96,15,1,Am),3),"128/2:",round(sim(128,15,2,Am),3))
# float32 case es=4
print("f32 C3-like (4,3,-5)")
Am=rot(4,3,-5)
for PX,xs in ((64,1),(64,2),(128,2),(96,2)):
    print(PX,xs,round(sim(PX,15,xs,Am,es=4),3))
