# ncu evidence for profiles/: launch list (gpu__time_duration) and one --set full capture of a whole step
# (the last of the three steps of `bench.py --profile-only`; 22 launches of the target render + 2 x 28 come first).
set -x
RX='regex:^(preprocess|seg_|scan_u32|duplicate_with|ranges_from|render_|ssim_|densify_)'
CMD="python bench.py --profile-only"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$RX" -s 78 -c 28 --csv --log-file gpurun_out/r1_launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k "$RX" -s 78 -c 28 -o gpurun_out/r1_prof $CMD > gpurun_out/ncu_f.log 2>&1
tail -3 gpurun_out/ncu_f.log
