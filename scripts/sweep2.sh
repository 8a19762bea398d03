python -m pytest tests/test_gpu_stream.py -m gpu -x -q 2>&1 | tail -3
python scripts/chain_latency.py 20000 | grep stream
show='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d["config"]["lanes_per_match"], {k: round(v,3) if v<1e6 else "%.3g"%v for k,v in d["phases"].items()}, round(d["roofline"]["frac"],4), "%.4g"%d["value"])'
for B in ${SB:-4}; do for L in ${LL:-1 2 4 8}; do echo "stream_blocks=$B lanes=$L"; TSRATE_STREAM_BLOCKS_PER_SM=$B timeout 300 python bench.py --path stream --steps 3 --warmup 2 --lanes $L --no-cpu-baseline --skip-e2e 2>&1 | python -c "$show"; done; done
