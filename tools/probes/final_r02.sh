# Final round-2 measurement pass on one B200: tests, launch lists, warm full captures, every bench line.
set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
B="python bench.py --no-e2e --no-cpu --no-latency --sub none --no-graph"
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_default_20.json 2> gpurun_out/f1.err; tail -2 gpurun_out/f1.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_reference_20.json 2>> gpurun_out/f1.err
python bench.py > gpurun_out/r02_bench_default_long.json 2>> gpurun_out/f1.err
python bench.py --workload c1 > gpurun_out/r02_bench_c1.json 2>> gpurun_out/f1.err
python bench.py --workload c5 > gpurun_out/r02_bench_c5.json 2>> gpurun_out/f1.err
SYMRELAY_EXTRACT_FUSED=0 python bench.py --workload c5 > gpurun_out/r02_bench_c5_two_kernels.json 2>> gpurun_out/f1.err
SYMRELAY_EXTRACT_FUSED=0 SYMRELAY_EXTRACT_OVERLAP=0 python bench.py --workload c5 > gpurun_out/r02_bench_c5_no_overlap.json 2>> gpurun_out/f1.err
python bench.py --workload c4sweep > gpurun_out/r02_bench_c4sweep.json 2>> gpurun_out/f1.err
python bench.py --workload c3 --no-latency --no-cpu > gpurun_out/r02_bench_c3.json 2>> gpurun_out/f1.err
BENCH_C3_SHARD_OF=8 python bench.py --workload c3 --no-latency --no-cpu > gpurun_out/r02_bench_c3_one_eighth.json 2>> gpurun_out/f1.err
SYMRELAY_WIDE=0 BENCH_C3_SHARD_OF=8 python bench.py --workload c3 --no-latency --no-cpu > gpurun_out/r02_bench_c3_one_eighth_no_wide.json 2>> gpurun_out/f1.err
SYMRELAY_PERSIST=0 python bench.py --steps 2000 --warmup 20 --sub none --no-latency --no-cpu > gpurun_out/r02_bench_c2_launch_per_batch.json 2>> gpurun_out/f1.err
python tools/latency_bench.py > gpurun_out/r02_latency_64k_streams.json 2>> gpurun_out/f1.err
tail -3 gpurun_out/f1.err
# launch lists (single pass, no replays): the kernel's share of a step
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 300 --csv --log-file gpurun_out/r02_c2_launches.csv $B --workload c2 --steps 200 --warmup 20 > gpurun_out/ncu_c2l.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 80 --csv --log-file gpurun_out/r02_c3_launches.csv $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3l.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 40 --csv --log-file gpurun_out/r02_c4_launches.csv $B --workload c4 --steps 6 --warmup 3 > gpurun_out/ncu_c4l.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 60 --csv --log-file gpurun_out/r02_c5_launches.csv $B --workload c5 --steps 10 --warmup 3 > gpurun_out/ncu_c5l.log 2>&1
BENCH_C3_SHARD_OF=8 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 80 --csv --log-file gpurun_out/r02_c3_one_eighth_launches.csv $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3el.log 2>&1
# full captures, warm (no cache flush between replays, many launches skipped)
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_packed_kernel --launch-skip 150 --launch-count 1 -o gpurun_out/r02_c2_packed $B --workload c2 --steps 200 --warmup 20 > gpurun_out/ncu_c2f.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_bulk_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c4_bulk $B --workload c4 --steps 6 --warmup 3 > gpurun_out/ncu_c4f.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_thread_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c3_thread $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3t.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_strided_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c3_strided $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3s.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_bulk_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c3_bulk $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3b.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_extract_seal_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c5_extract_seal $B --workload c5 --steps 10 --warmup 3 > gpurun_out/ncu_c5f.log 2>&1
SYMRELAY_EXTRACT_FUSED=0 SYMRELAY_EXTRACT_OVERLAP=0 ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_extract_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c5_extract $B --workload c5 --steps 10 --warmup 3 > gpurun_out/ncu_c5e.log 2>&1
SYMRELAY_EXTRACT_FUSED=0 SYMRELAY_EXTRACT_OVERLAP=0 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 60 --csv --log-file gpurun_out/r02_c5_two_kernels_launches.csv $B --workload c5 --steps 10 --warmup 3 > gpurun_out/ncu_c5l2.log 2>&1
BENCH_C3_SHARD_OF=8 ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_wide_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c3_one_eighth_wide $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3w.log 2>&1
for f in r02_c2_packed r02_c4_bulk r02_c3_thread r02_c3_strided r02_c3_bulk r02_c5_extract r02_c5_extract_seal r02_c3_one_eighth_wide; do
  ncu -i gpurun_out/$f.ncu-rep --page details > gpurun_out/${f}_ncu_full.txt 2>&1
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/${f}_raw.csv 2>&1
done
ncu -i gpurun_out/r02_c2_packed.ncu-rep --page source --csv --print-source sass > gpurun_out/r02_c2_packed_source.csv 2>&1
ls -la gpurun_out/*.ncu-rep
python -c "from symmetry_b200 import build; build.build(force=True, extra=['-DSS_PROBE_TIMELINE'], out='/root/repo/symmetry_b200/libsymrelay_probe.so')" > /dev/null 2>&1
SYMRELAY_SO=symmetry_b200/libsymrelay_probe.so python tools/probes/c5_timeline.py > gpurun_out/r02_c5_timeline_fused.txt 2>&1
SYMRELAY_EXTRACT_FUSED=0 SYMRELAY_SO=symmetry_b200/libsymrelay_probe.so python tools/probes/c5_timeline.py > gpurun_out/r02_c5_timeline_two_kernels.txt 2>&1
rm -f symmetry_b200/libsymrelay_probe.so
