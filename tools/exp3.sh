mkdir -p gpurun_out
line() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'ms_per_step', round(d['ms_per_step'],4), 'value', '%.4g' % d['value'], 'e2e', '%.4g' % (d.get('e2e') or {}).get('value', 0))"; }
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
timeout 200 python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>&1 | line "G9n4"
DRL_FUSE_KEYS=1 timeout 200 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-e2e 2>&1 | line "G9n4 fused keys"
timeout 200 python bench.py --grid 8 --drones 3 --steps 40 --warmup 5 --no-cpu-baseline 2>&1 | line "G8n3"
