"""One tcgen05 dense layer at the DeformationNetwork's size (for ncu captures): python tools/gemm_one.py [rows K N]"""
import sys, os, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "animating-gaussian-splats_b200")]
from gsb200.linear import linear
rows, K, N = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (300000, 192, 256)
x = torch.randn(rows, K, device="cuda"); w = torch.randn(N, K, device="cuda"); b = torch.randn(N, device="cuda")
y = torch.empty(rows, N, device="cuda")
for _ in range(3):
    linear(x, w, b, out=y)
torch.cuda.synchronize()
