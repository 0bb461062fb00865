import sys, numpy as np, torch
sys.path[:0]=['/root/repo','/root/repo/animating-gaussian-splats_b200','/root/repo/tests']
from gsb200 import rigidity as R
from oracle import rigidity_ref
dev=torch.device('cuda'); gen=torch.Generator().manual_seed(5)
P,K=300_000,20
means=(torch.rand(P,3,generator=gen)-0.5)*3; quats=torch.randn(P,4,generator=gen)
seg=torch.zeros(P,3); seg[:,0]=(torch.rand(P,generator=gen)>0.4).float(); seg[:,2]=1-seg[:,0]
F=int(seg[:,0].sum())
walk=torch.cumsum(0.02*torch.randn(F,3,generator=gen),0); means[seg[:,0]>0.5]=walk-walk.mean(0)
steps=torch.tensor([d for d in range(-10,11) if d!=0]); nbr=(torch.arange(F)[:,None]+steps[None])%F
w=torch.rand(F,K,generator=gen)
base={"means":means.to(dev),"rotation_quaternions":quats.to(dev),"segmentation_masks":seg.to(dev)}
ninfo=R.NeighborInfo(weights=w.to(dev),indices=nbr.to(dev))
prev=R.create_foreground_info(base,ninfo.indices)
jit={"means":(means+0.01*torch.randn(P,3,generator=gen)).to(dev).requires_grad_(True),"rotation_quaternions":quats.to(dev).requires_grad_(True),"segmentation_masks":base["segmentation_masks"]}
import time
loss=R.calculate_rigidity_loss(jit,ninfo,prev); loss.backward()
fg=np.nonzero(seg[:,0].numpy()>0.5)[0]
a=(jit["means"].detach().cpu().numpy(),quats.numpy(),fg,nbr.numpy(),w.numpy(),prev.inverted_rotations.cpu().numpy(),prev.offsets_to_neighbors.cpu().numpy())
want,dm,dq=rigidity_ref.rigidity_loss(*a)
w32,dm32,dq32=rigidity_ref.rigidity_loss(*a,dtype=np.float32)
got=jit["rotation_quaternions"].grad.cpu().numpy().astype(np.float64)
tol=1e-4*np.abs(dq)+2e-5*np.abs(dq).max(axis=1,keepdims=True)+1e-12
bad=np.argwhere(np.abs(got-dq)>tol)
print("bad",len(bad))
for r,c in bad[:10]:
    print(r,c,"|q|",np.linalg.norm(quats[r].numpy()),"got",got[r],"want",dq[r],"f32",dq32[r])
# timing
torch.cuda.synchronize()
for grad in (True,False):
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    for it in range(3):
        e0.record()
        for _ in range(20):
            if grad:
                l=R.calculate_rigidity_loss(jit,ninfo,prev)
            else:
                with torch.no_grad(): l=R.calculate_rigidity_loss(jit,ninfo,prev)
        e1.record(); torch.cuda.synchronize()
    print("grad",grad,"ms/call",e0.elapsed_time(e1)/20, "F",F)
e0.record()
for _ in range(20): R.create_foreground_info(base,ninfo.indices)
e1.record(); torch.cuda.synchronize(); print("fginfo ms",e0.elapsed_time(e1)/20)
