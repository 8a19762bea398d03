# Round-2 profile set (run under gpurun): launch list of one default bench step (both paths), full
# captures of the dominant kernels at config 2, then the rating kernels on ladders whose rows do not
# fit L2 (prof_big.sh).  scripts/make_profiles.py turns the reports into profiles/r2_*.
set -x
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --skip-e2e --no-configs"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_rate_stream|k_rate<|k_radix_onesweep|k_pred|k_radix_hist_matches|k_levels' -c 16 -f -o gpurun_out/r2_prof $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log | cut -c1-200
bash scripts/prof_big.sh
