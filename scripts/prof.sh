# launch list + full capture of selected kernels (run under gpurun)
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --skip-e2e --lanes ${LANES:-2} ${BENCH_ARGS:-}"
export TSRATE_BLOCKS_PER_SM=${TSRATE_BLOCKS_PER_SM:-1}
KREGEX=${KREGEX:-k_rate|k_frontier}
NK=${NK:-2}
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$KREGEX" -c $NK -o gpurun_out/prof $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log | cut -c1-200
