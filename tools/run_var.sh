p() { python -c "import json,sys; d=json.loads(sys.stdin.read()); s=d['stage_ms']; print(round(d['ms_per_step'],3), {k: round(v,3) for k,v in s.items() if k in ('depth_sort','sort','duplicate_keys','scan','tile_ranges')})"; }
timeout 300 python -m pytest tests/test_binning_gpu.py tests/test_raster_gpu.py -m gpu -x -q 2>&1 | tail -3
GSB_SS_OPT=0 timeout 300 python -m pytest tests/test_binning_gpu.py -m gpu -x -q 2>&1 | tail -3
echo OPT1; python bench.py --no-cpu-baseline --no-reference-cuda 2>/dev/null | p
echo OPT0; GSB_SS_OPT=0 python bench.py --no-cpu-baseline --no-reference-cuda 2>/dev/null | p
echo D11; GSB_DEPTH_DIGIT=11 python bench.py --no-cpu-baseline --no-reference-cuda 2>/dev/null | p
