nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
echo "== single process 8"; python bench.py --single-process --gpus 8 --steps 2000 --warmup 20 > gpurun_out/sp8.json 2> gpurun_out/sp8.err; tail -2 gpurun_out/sp8.err
echo "== single process 8, 20 steps"; python bench.py --single-process --gpus 8 --steps 20 --warmup 5 > gpurun_out/sp8s.json 2>> gpurun_out/sp8.err
echo "== torchrun 8, 20 steps"; SYMRELAY_PROF=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/tr8s.json 2> gpurun_out/tr8s.err; grep profile gpurun_out/tr8s.err | head -8
echo "== torchrun 8, 2000 steps"; SYMRELAY_PROF=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 2000 --warmup 20 --sub none > gpurun_out/tr8.json 2> gpurun_out/tr8.err; grep profile gpurun_out/tr8.err | head -8
echo "== reference 8"; python bench.py --impl reference --gpus 8 --steps 20 --warmup 5 > gpurun_out/ref8.json 2> gpurun_out/ref8.err
