timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
python - <<'P'
import sys, json
sys.path[:0]=['/root/repo','/root/repo/animating-gaussian-splats_b200','/root/repo/tests']
import bench
class A: gaussians=300000; cams=27; width=640; height=360
print("densify-shaped ours:", bench.densify_shaped(A, "ours"))
P
python bench.py --no-cpu-baseline --no-reference-cuda --steps 20 2>/dev/null | python -c "import sys,json; d=json.load(sys.stdin); print(round(d['ms_per_step'],3), d['e2e'])"
