CMD="python bench.py --matches 125000000 --players 12500000 --steps 1 --warmup 1 --no-cpu-baseline --skip-e2e"
$CMD > gpurun_out/plain_big.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_rate_stream|k_rate<' -c 2 -o gpurun_out/prof_big $CMD > gpurun_out/ncu_big.log 2>&1
tail -2 gpurun_out/ncu_big.log | cut -c1-200
