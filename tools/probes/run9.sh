timeout 800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --workload c3 --no-e2e --no-cpu --no-latency 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('c3', d['ms_per_step']*1e3, d['alu_roofline']['frac'])"
python bench.py --workload c4 --no-e2e --no-cpu --no-latency 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('c4', d['ms_per_step']*1e3, d['alu_roofline']['frac'])"
python bench.py --workload c2 --no-e2e --no-cpu --no-latency --sub none 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('c2', d['ms_per_step']*1e3, d['alu_roofline']['frac'])"
