set -x
timeout 800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
B="python bench.py --no-e2e --no-cpu --no-latency --sub none --no-graph"
# launch lists (single pass, no replays): the kernel's share of a step
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 300 --csv --log-file gpurun_out/r02_c2_launches.csv $B --workload c2 --steps 200 --warmup 20 > gpurun_out/ncu_c2l.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 80 --csv --log-file gpurun_out/r02_c3_launches.csv $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3l.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 40 --csv --log-file gpurun_out/r02_c4_launches.csv $B --workload c4 --steps 6 --warmup 3 > gpurun_out/ncu_c4l.log 2>&1
# full captures, warm (no cache flush between replays, many launches skipped)
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_packed_kernel --launch-skip 150 --launch-count 1 -o gpurun_out/r02_c2_packed $B --workload c2 --steps 200 --warmup 20 > gpurun_out/ncu_c2f.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_bulk_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c4_bulk $B --workload c4 --steps 6 --warmup 3 > gpurun_out/ncu_c4f.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_thread_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c3_thread $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3t.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_strided_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c3_strided $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3s.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:ss_bulk_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_c3_bulk $B --workload c3 --steps 10 --warmup 3 > gpurun_out/ncu_c3b.log 2>&1
for f in r02_c2_packed r02_c4_bulk r02_c3_thread r02_c3_strided r02_c3_bulk; do
  ncu -i gpurun_out/$f.ncu-rep --page details > gpurun_out/${f}_ncu_full.txt 2>&1
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/${f}_raw.csv 2>&1
done
ls -la gpurun_out/*.ncu-rep
