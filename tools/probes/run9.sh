for rep in 1 2 3; do
BENCH_SYNC_LAT=0 timeout 300 python bench.py --steps 20 --warmup 5 --sub none --no-latency --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('20 steps', d['e2e']['value']/1e6, d['e2e']['ms_per_step']*1e3)"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 1 --steps 20 --warmup 5 --sub none --no-latency --no-cpu 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('torchrun 1', d['e2e']['value']/1e6, d['e2e']['ms_per_step']*1e3)"
