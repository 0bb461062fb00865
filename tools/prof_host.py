"""cProfile of the densify-shaped single-camera loop (host overhead of the drop-in path), sorted by own time."""
import cProfile, pstats, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
class A: gaussians=300000; cams=27; width=640; height=360
bench.densify_shaped(A, "ours", iters=5, warm=3)
pr = cProfile.Profile(); pr.enable()
r = bench.densify_shaped(A, "ours", iters=200, warm=3)
pr.disable()
print("iters/s", r)
pstats.Stats(pr).sort_stats("tottime").print_stats(32)
