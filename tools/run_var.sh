python - <<'P'
import sys, json
sys.path[:0]=['/root/repo','/root/repo/animating-gaussian-splats_b200','/root/repo/tests']
import bench
class A: gaussians=300000; cams=27; width=640; height=360
print(json.dumps(bench.deformation_shaped(A), indent=1))
P
