"""Dev tool (CPU only): per-z-layer tile-class counts of cfg4 (numpy restatement of
fuse_classify_kernel on the oracle's 2D SDF) saved to /tmp/per_layer.npy; the weights of the
auto-balance cost model (fuse_layer_cost_kernel) were fitted to these counts and the fusion times
of 8 equal slabs (profiles/r02c_bench_n8_equal_slabs.json)."""
import numpy as np, sys, time
sys.path.insert(0, '.')
from oracle import bindings as B
from vacancy_b200 import synthetic as S
port=B.PortLib()
n=1024; size=2048
cams=S.circle_cameras(360,size)
pos=((np.arange(n)+0.5)*(2/n)-1).astype(np.float32)
sdf=port.make_sdf(S.silhouette(cams[0]), None, True, True, 0.1)
ones=(sdf==1.0).reshape(size//8,8,size//8,8).all(axis=(1,3))
kval=(sdf==sdf.min()).reshape(size//8,8,size//8,8).all(axis=(1,3))
def integ(a):
    I=np.zeros((a.shape[0]+1,a.shape[1]+1),np.int64); I[1:,1:]=np.cumsum(np.cumsum(a,0),1); return I
Io, Ik = integ(ones), integ(kval)
tx_sz,ty_sz,tz_sz=32,8,8
X0=pos[::tx_sz]; X1=pos[tx_sz-1::tx_sz]; Y0=pos[::ty_sz]; Y1=pos[ty_sz-1::ty_sz]; Z0=pos[::tz_sz]; Z1=pos[tz_sz-1::tz_sz]
shape=(len(Z0),len(Y0),len(X0))
per_layer=np.zeros((360,len(Z0),5))
t=time.time()
for vi in range(0,360,3):
    c=cams[vi]
    w2c=np.linalg.inv(c.c2w()); f=1.2*size; cc=size/2-0.5
    umin=np.full(shape,1e30); umax=-umin.copy(); vmin=umin.copy(); vmax=-umin.copy()
    for zs in (Z0,Z1):
        for ys in (Y0,Y1):
            for xs in (X0,X1):
                X=xs[None,None,:]; Y=ys[None,:,None]; Z=zs[:,None,None]
                pc=[w2c[i,0]*X+w2c[i,1]*Y+w2c[i,2]*Z+w2c[i,3] for i in range(3)]
                u=f*pc[0]/pc[2]+cc; v=f*pc[1]/pc[2]+cc
                umin=np.minimum(umin,u);umax=np.maximum(umax,u);vmin=np.minimum(vmin,v);vmax=np.maximum(vmax,v)
    ua=np.floor(umin).astype(np.int64)-1; ub=np.floor(umax).astype(np.int64)+2
    va=np.floor(vmin).astype(np.int64)-1; vb=np.floor(vmax).astype(np.int64)+2
    inside=(ua>=0)&(ub<=size-1)&(va>=0)&(vb<=size-1)
    outside=(~inside)&((umax+2<0)|(umin-2>size-1)|(vmax+2<0)|(vmin-2>size-1))
    bx0=np.clip(ua//8,0,size//8-1); bx1=np.clip(ub//8,0,size//8-1); by0=np.clip(va//8,0,size//8-1); by1=np.clip(vb//8,0,size//8-1)
    area=(bx1-bx0+1)*(by1-by0+1)
    def rectsum(I): return I[by1+1,bx1+1]-I[by0,bx1+1]-I[by1+1,bx0]+I[by0,bx0]
    allones=inside&(rectsum(Io)==area); allk=inside&(rectsum(Ik)==area)
    cls=np.zeros(shape,int); cls[inside]=3; cls[outside]=4; cls[allones]=1; cls[allk]=2
    for k in range(5): per_layer[vi,:,k]=(cls==k).sum(axis=(1,2))
print('time',time.time()-t)
np.save('/tmp/per_layer.npy', per_layer)
