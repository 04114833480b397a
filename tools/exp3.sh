mkdir -p gpurun_out
line() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],4), 'value', '%.4g' % d['value'], 'e2e', '%.4g' % (d.get('e2e') or {}).get('value', 0))"; }
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q -m gpu 2>&1 | tail -2
timeout 200 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | line "fused-keys"
DRL_NO_FUSE=1 timeout 200 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | line "nofuse"
timeout 200 python bench.py --grid 8 --drones 3 --steps 200 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | line "G8n3"
