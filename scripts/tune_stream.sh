show='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d["phases"]["rate_ms"])'
for NS in 0 20 50 100; do echo -n "sleep_ns=$NS rate_ms="; TSRATE_STREAM_SLEEP_NS=$NS timeout 120 python bench.py --path stream --steps 3 --warmup 2 --no-cpu-baseline --skip-e2e 2>&1 | python -c "$show"; done
cp analyzer_b200/libtsrate.so /tmp/orig.so; cp analyzer_b200/libtsrate_mb3.so analyzer_b200/libtsrate.so
for B in 3; do echo -n "min_blocks=3 blocks/SM=$B rate_ms="; TSRATE_STREAM_BLOCKS_PER_SM=$B timeout 120 python bench.py --path stream --steps 3 --warmup 2 --no-cpu-baseline --skip-e2e 2>&1 | python -c "$show"; done
cp /tmp/orig.so analyzer_b200/libtsrate.so
