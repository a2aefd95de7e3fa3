set -x
B="python bench.py --no-e2e --no-cpu --no-latency --sub none --no-graph"
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_default_20.json 2> gpurun_out/f1.err
python bench.py > gpurun_out/r02_bench_default_long.json 2>> gpurun_out/f1.err
python bench.py --workload c3 --no-latency --no-cpu > gpurun_out/r02_bench_c3.json 2>> gpurun_out/f1.err
BENCH_C3_SHARD_OF=8 python bench.py --workload c3 --no-latency --no-cpu > gpurun_out/r02_bench_c3_one_eighth.json 2>> gpurun_out/f1.err
python bench.py --workload c4sweep > gpurun_out/r02_bench_c4sweep.json 2>> gpurun_out/f1.err
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 80 --csv --log-file gpurun_out/r02_c3_launches.csv $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3l.log 2>&1
python -c "from symmetry_b200 import build; build.build(force=True, extra=['-DSS_PROBE_TIMELINE'], out='/root/repo/symmetry_b200/libsymrelay_probe.so')" > /dev/null 2>&1
SYMRELAY_SO=symmetry_b200/libsymrelay_probe.so python tools/probes/c3_timeline.py > gpurun_out/r02_c3_timeline.txt 2>&1
rm -f symmetry_b200/libsymrelay_probe.so
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
tail -2 gpurun_out/f1.err
