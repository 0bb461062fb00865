"""cProfile of the train.py-shaped optimiser step at a tiny size (host overhead of the per-timestep path), by own time."""
import cProfile, pstats, sys, os, types, io
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "animating-gaussian-splats_b200"))
import bench
a = types.SimpleNamespace(gaussians=3000, cams=27, width=64, height=48, seed=0)
bench.train_step_shaped(a, iters=4)
pr = cProfile.Profile(); pr.enable()
r = bench.train_step_shaped(a, iters=60)
pr.disable()
print({k: r[k] for k in ("ours", "ours_tf32", "stock")})
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(45)
print(s.getvalue()[:9000])
