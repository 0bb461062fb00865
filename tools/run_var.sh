# scratch: tests + the extra bench legs
timeout 800 python -m pytest tests/test_rigidity.py -x -q -m gpu 2>&1 | tail -5
python - <<'P'
import sys, json
sys.path[:0]=['/root/repo','/root/repo/animating-gaussian-splats_b200','/root/repo/tests']
import bench
class A: gaussians=300000; cams=27; width=640; height=360
print(json.dumps(bench.rigidity_shaped(A), indent=1))
P
