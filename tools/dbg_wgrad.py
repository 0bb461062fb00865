import sys, torch
sys.path[:0]=['/root/repo','/root/repo/animating-gaussian-splats_b200']
from gsb200.linear import linear_wgrad
rows,N,K=32,128,32
for (r0,n0) in [(0,0),(1,0),(0,1),(0,4),(0,33),(8,0),(9,5)]:
    G=torch.zeros(rows,N,device='cuda'); G[r0,n0]=1
    X=(100*torch.arange(rows,device='cuda')[:,None]+torch.arange(K,device='cuda')[None]).float()
    dw=linear_wgrad(G,X); torch.cuda.synchronize()
    nz=dw.nonzero()
    print((r0,n0), "nonzero rows:", sorted(set(nz[:,0].tolist()))[:8], "count", len(nz))
    if len(nz):
        n1=int(nz[0,0]); print("   row", n1, dw[n1,:12].tolist())
