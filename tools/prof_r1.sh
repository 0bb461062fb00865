set -x
CMD="python bench.py --profile-only"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 120 -c 60 --csv --log-file gpurun_out/r1_launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"rs_onesweep|rs_histogram|render_bwd|render_fwd|duplicate_with|tile_ranges|ssim_" -s 40 -c 14 -o gpurun_out/r1_prof $CMD > gpurun_out/ncu_f.log 2>&1
tail -3 gpurun_out/ncu_f.log
ls -la gpurun_out
