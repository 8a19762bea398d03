python -m pytest tests -m gpu -x -q 2>&1 | tail -3
show='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d["config"]["lanes_per_match"], {k: round(v,3) if v<1e6 else "%.3g"%v for k,v in d["phases"].items()}, round(d["roofline"]["frac"],4), "%.4g"%d["value"])'
for B in ${RB:-1 2}; do for L in ${LL:-2 4 8}; do for F in ${FB:-2}; do echo "rate_blocks=$B lanes=$L frontier_blocks=$F"; TSRATE_FRONTIER_BLOCKS_PER_SM=$F TSRATE_BLOCKS_PER_SM=$B python bench.py --steps 3 --warmup 2 --lanes $L --no-cpu-baseline --skip-e2e 2>&1 | python -c "$show"; done; done; done
