timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for d in 12; do
echo "== depth $d"
BENCH_E2E_DEPTH=$d BENCH_SYNC_LAT=0 SYMRELAY_TS=1 SYMRELAY_PROF=1 timeout 300 python bench.py --steps 3000 --warmup 20 --sub none --no-latency --no-cpu 2>gpurun_out/b19_$d.err >gpurun_out/b19_$d.json; grep -E "resident|profile|latest|chain" gpurun_out/b19_$d.err
done
echo "== sync"
SYMRELAY_TS=1 SYMRELAY_PROF=1 timeout 300 python bench.py --steps 20 --warmup 5 --sub none --no-latency --no-cpu 2>gpurun_out/b19s.err >gpurun_out/b19s.json; grep -E "resident|profile|latest|chain" gpurun_out/b19s.err
