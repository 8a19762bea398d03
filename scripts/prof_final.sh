# Round-1 profile set (run under gpurun): launch list of one default bench step (both paths),
# then full captures of the dominant kernels of each path.
set -x
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --skip-e2e"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_rate_stream|k_rate<|k_frontier|k_radix_onesweep|k_pred|k_link|k_radix_hist_all|k_radix_hist_matches' -c 24 -o gpurun_out/prof $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log | cut -c1-200
