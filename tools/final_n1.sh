set -x
mkdir -p gpurun_out
(time python -m pytest tests -x -q -m gpu) > gpurun_out/gputests.log 2>&1; tail -3 gpurun_out/gputests.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
(time python bench.py) > gpurun_out/bench_n1.log 2> gpurun_out/bench_n1.err; tail -c 600 gpurun_out/bench_n1.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/cpu_arm_n1.log 2>/dev/null; tail -c 400 gpurun_out/cpu_arm_n1.log
export GSB_BENCH_EXTRAS=0
OURS="python bench.py --profile-only --no-cpu-baseline --no-reference-cuda --no-c5"
ncu --metrics gpu__time_duration.sum --clock-control none --graph-profiling node -c 700 --csv --log-file gpurun_out/r2_launches.csv $OURS > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on --graph-profiling node -k regex:'render_|preprocess_|seg_|scan_u32|duplicate_with|ssim_|flat_|cam_params|ranges_|tile_' -s 64 -c 34 -o gpurun_out/r2_prof $OURS > gpurun_out/ncu_f.log 2>&1
tail -2 gpurun_out/ncu_f.log
ls -la gpurun_out/r2_prof.ncu-rep
