# scratch: GEMM timing against cuBLAS
python - <<'P'
import sys, torch
sys.path[:0]=['/root/repo','/root/repo/animating-gaussian-splats_b200']
from gsb200.linear import linear
def t(fn, n=20):
    for _ in range(3): fn()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1)/n
for rows,K,N in ((300000,192,256),(300000,256,256),(300000,128,128),(300000,256,7),(1000000,192,256)):
    x=torch.randn(rows,K,device='cuda'); w=torch.randn(N,K,device='cuda'); b=torch.randn(N,device='cuda')
    y=torch.empty(rows,N,device='cuda')
    ours=t(lambda: linear(x,w,b,out=y)); ours3=t(lambda: linear(x,w,b,precise=True,out=y))
    torch.backends.cuda.matmul.allow_tf32=False
    f32=t(lambda: torch.addmm(b,x,w.t(),out=y))
    torch.backends.cuda.matmul.allow_tf32=True
    tf32=t(lambda: torch.addmm(b,x,w.t(),out=y))
    byts=rows*(K+N)*4
    print(f"{rows}x{K}x{N}: ours {ours*1e3:.0f} us ({byts/ours/1e6:.0f} GB/s, {2*rows*K*N/ours/1e9:.0f} TF/s)  ours3x {ours3*1e3:.0f} us  cublas fp32 {f32*1e3:.0f} us  cublas tf32 {tf32*1e3:.0f} us")
P
