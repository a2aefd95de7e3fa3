for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r02_bench_c2_n${n}_20steps.json 2> gpurun_out/tr${n}s.err
  tail -1 gpurun_out/tr${n}s.err
done
