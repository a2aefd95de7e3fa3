timeout 800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for k in 20 3000; do
BENCH_SYNC_LAT=$([ $k = 20 ] && echo 1 || echo 0) timeout 300 python bench.py --steps $k --warmup 5 --sub none --no-latency --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('steps', d['steps'], 'e2e', d['e2e']['value']/1e6, d['e2e']['ms_per_step']*1e3, d['e2e'].get('batch_latency',{}).get('p50_us'))"
done
BENCH_SWEEP_SIZES=2048,4096,8192,16384 python bench.py --workload c4sweep 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print([(p['frame_bytes'], p['streams'], round(p['alu_frac'],3), p['kernel']) for p in d['sweep']])"
