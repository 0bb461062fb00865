timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for i in 1 2; do python bench.py --steps 10 --no-cpu-baseline --no-reference-cuda 2>gpurun_out/e.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); s=d['stage_ms']; print(round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3), 'raster stages', round(sum(s.values()),3), {k: round(v,3) for k,v in s.items()})"; done
