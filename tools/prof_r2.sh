# Round-2 ncu evidence (run on the GPU box from the repo root): launch lists of our step and of the unmodified
# reference extension on the same workload, one --set full capture of our step, all under gpurun_out/.
set -x
export GSB_BENCH_EXTRAS=0
RX='regex:^(gsb::|void gsb::)'
OURS="python bench.py --profile-only --no-cpu-baseline --no-reference-cuda --no-c5"
$OURS > gpurun_out/plain.log 2>&1 || exit 1
# eager steps of --profile-only come first (3 x one_step after the parity step), then graph replays
ncu --metrics gpu__time_duration.sum --clock-control none --graph-profiling node -c 700 --csv --log-file gpurun_out/r2_launches.csv $OURS > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on --graph-profiling node -k regex:'render_|preprocess_|seg_|scan_u32|duplicate_with|ssim_|flat_|cam_params|ranges_|tile_' -s 64 -c 34 -o gpurun_out/r2_prof $OURS > gpurun_out/ncu_f.log 2>&1
REF="python bench.py --impl reference_cuda --steps 1"
GSB_REF_PROFILE=1 $REF > gpurun_out/ref_plain.log 2>&1 || exit 1
GSB_REF_PROFILE=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r2_ref_launches.csv $REF > gpurun_out/ncu_r.log 2>&1
tail -3 gpurun_out/ncu_f.log gpurun_out/ncu_r.log
