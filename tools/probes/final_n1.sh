python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_default_20.json 2> gpurun_out/f1.err; tail -2 gpurun_out/f1.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_reference_20.json 2>> gpurun_out/f1.err
python bench.py > gpurun_out/r02_bench_default_long.json 2>> gpurun_out/f1.err
python bench.py --workload c1 > gpurun_out/r02_bench_c1.json 2>> gpurun_out/f1.err
python bench.py --workload c5 > gpurun_out/r02_bench_c5.json 2>> gpurun_out/f1.err
python bench.py --workload c4sweep > gpurun_out/r02_bench_c4sweep.json 2>> gpurun_out/f1.err
python bench.py --workload c3 --no-latency --no-cpu > gpurun_out/r02_bench_c3.json 2>> gpurun_out/f1.err
SYMRELAY_PERSIST=0 python bench.py --steps 2000 --warmup 20 --sub none --no-latency --no-cpu > gpurun_out/r02_bench_c2_launch_per_batch.json 2>> gpurun_out/f1.err
python tools/latency_bench.py > gpurun_out/r02_latency_64k_streams.json 2>> gpurun_out/f1.err
tail -3 gpurun_out/f1.err
