# Final round-2 multi-GPU pass (run with gpurun --gpus 8): the driver's scaling form at N = 2, 4, 8, the long form at 8,
# and one process driving all eight.
set -x
nvidia-smi topo -m > gpurun_out/r02_topo_8gpu.txt 2>&1
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r02_bench_c2_n${n}_20steps.json 2> gpurun_out/tr${n}s.err
  tail -2 gpurun_out/tr${n}s.err
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 2000 --warmup 20 > gpurun_out/r02_bench_c2_n8_2000steps.json 2> gpurun_out/tr8.err
tail -2 gpurun_out/tr8.err
python bench.py --single-process --gpus 8 --steps 2000 --warmup 20 > gpurun_out/r02_bench_c2_single_process_n8.json 2> gpurun_out/sp8.err
tail -2 gpurun_out/sp8.err
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
