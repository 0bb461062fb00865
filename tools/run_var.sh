p() { python -c "import json,sys; d=json.loads(sys.stdin.read()); s=d['stage_ms']; print(round(d['ms_per_step'],3), {k: round(v,3) for k,v in s.items() if k in ('render_fwd','render_bwd')})"; }
echo base; python bench.py --steps 5 --no-cpu-baseline --no-reference-cuda 2>/dev/null | p
for v in 3_5 5_7 6_8 2_4; do echo $v; GSB200_LIB=$PWD/animating-gaussian-splats_b200/csrc/var/lib_$v.so python bench.py --steps 5 --no-cpu-baseline --no-reference-cuda 2>/dev/null | p; done
