"""cProfile of gsb200.timestep_train.train_step at a tiny size (GPU time negligible: what is left is the host's issue time
per optimiser step), by own time and by cumulative time."""
import cProfile, io, os, pstats, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "animating-gaussian-splats_b200"))
import time
import torch
from gsb200 import deform, rigidity, scene, timestep_train as tt

dev = torch.device("cuda")
P, W, H, V = 3000, 64, 48, 5
raw = {k: v.to(dev) for k, v in scene.make_gaussians(P, seed=0, extent=1.0).items()}
stacked = scene.stack_cameras(scene.make_cameras(V, W, H, seed=0), dev)
targets = torch.rand(V, 3, H, W, device=dev)
ninfo = rigidity.create_neighbor_info(raw, k=20)
enc0 = deform.normalize_and_encode_means_and_rotations(raw)
fg0 = rigidity.create_foreground_info(raw, ninfo.indices)
net = deform.DeformationNetwork(256, 4, precise=True).to(dev)
opt = torch.optim.Adam(net.parameters(), lr=1e-4)
va = tt.ViewArena()
step = lambda: tt.train_step(net, opt, raw, enc0, enc0, fg0, ninfo, (stacked, targets), 7, 150, view_arena=va)
for _ in range(5):
    step()
torch.cuda.synchronize()
n = 100
t0 = time.perf_counter()
pr = cProfile.Profile(); pr.enable()
for _ in range(n):
    step()
pr.disable()
torch.cuda.synchronize()
print(f"host ms per step (under cProfile): {(time.perf_counter() - t0) / n * 1e3:.3f}")
t0 = time.perf_counter()
for _ in range(n):
    step()
torch.cuda.synchronize()
print(f"ms per step (no profiler): {(time.perf_counter() - t0) / n * 1e3:.3f}")
for key, k in (("tottime", 28), ("cumulative", 40)):
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats(key).print_stats(k)
    print("\n".join(l[:150] for l in s.getvalue().splitlines()[6:]))
