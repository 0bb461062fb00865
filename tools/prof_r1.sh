# ncu evidence for profiles/: launch list (gpu__time_duration) and one --set full capture of a whole step.
set -x
CMD="python bench.py --profile-only"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^(preprocess|rs_|scan_u32|duplicate_with|tile_ranges|render_|ssim_|densify_)" -s 21 -c 21 --csv --log-file gpurun_out/r1_launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"^(preprocess|rs_|scan_u32|duplicate_with|tile_ranges|render_|ssim_|densify_)" -s 21 -c 21 -o gpurun_out/r1_prof $CMD > gpurun_out/ncu_f.log 2>&1
tail -3 gpurun_out/ncu_f.log
