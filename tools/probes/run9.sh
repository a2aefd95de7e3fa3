for c in "1 4" "2 4" "3 4" "2 6" "3 8"; do set -- $c
for k in 20 3000; do
echo "== ctas/sm $1 drain_every $2 steps $k"
SYMRELAY_PERSIST_CTAS_PER_SM=$1 SYMRELAY_PERSIST_DRAIN_EVERY=$2 BENCH_SYNC_LAT=0 timeout 300 python bench.py --steps $k --warmup 5 --sub none --no-latency --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['e2e']['value']/1e6, d['e2e']['ms_per_step']*1e3)"
done; done
