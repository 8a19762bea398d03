#!/bin/bash
# Rate-phase time of both paths for every lanes-per-match setting (DESIGN.md 3.2's table) and the
# poll back-off of the streaming kernel.  Run under gpurun; prints one line per setting.
cd "$(dirname "$0")/.."
one() {
    timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --skip-e2e "$@" 2>&1 | tail -1 | python -c "
import json, sys
d = json.loads(sys.stdin.read())
p = d['paths']
print('  stream link %.2f rate %.2f | rounds schedule %.2f rate %.2f ms' % (
    p['stream']['schedule_ms'], p['stream']['rate_ms'], p['rounds']['schedule_ms'], p['rounds']['rate_ms']))"
}
for l in 1 2 4 8; do echo "3v3 config 2, lanes=$l"; one --lanes $l; done
for l in 2 4 8 16; do echo "5v5 10M matches 1M players Zipf, lanes=$l"; one --slots 10 --zipf --p-draw 0.02 --lanes $l; done
for l in 2 4 8; do echo "5v5 6M matches 1M players uniform, lanes=$l"; one --slots 10 --matches 6000000 --lanes $l; done
for ns in 0 20 100 400; do echo "3v3 config 2, TSRATE_STREAM_SLEEP_NS=$ns"; TSRATE_STREAM_SLEEP_NS=$ns one; done
