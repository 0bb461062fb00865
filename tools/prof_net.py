"""DeformationNetwork forward + backward (300 k Gaussians, 256 x 4 blocks, precise mode) for an ncu launch list:
   ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/net_launches.csv python tools/prof_net.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "animating-gaussian-splats_b200"))
from gsb200 import deform

n = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000
torch.manual_seed(0)
net = deform.DeformationNetwork(256, 4, precise=True).cuda()
x = torch.randn(n, 192, device="cuda")
init = torch.randn(n, 7, device="cuda")
w = torch.randn(n, 7, device="cuda")
for it in range(3):
    net.zero_grad(set_to_none=True)
    torch.cuda.synchronize()
    (net(init, network_input=x) * w).sum().backward()
    torch.cuda.synchronize()
print("ok")
