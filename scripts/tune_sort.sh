show='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d["phases"]["schedule_ms"])'
echo -n "default link_ms="; timeout 120 python bench.py --path stream --steps 5 --warmup 2 --no-cpu-baseline --skip-e2e 2>&1 | python -c "$show"
cp analyzer_b200/libtsrate.so /tmp/orig.so
for V in i8b8 i8b6 i12b6; do cp analyzer_b200/libtsrate_$V.so analyzer_b200/libtsrate.so; echo -n "$V link_ms="; timeout 120 python bench.py --path stream --steps 5 --warmup 2 --no-cpu-baseline --skip-e2e 2>&1 | python -c "$show"; done
cp /tmp/orig.so analyzer_b200/libtsrate.so
