#!/bin/bash
# C4: warp-specialised fast kernel (default) vs the sequential one (DRL_FAST_SPLIT=0); parity first
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_gpu_guard_bands.py tests/test_gpu_full_size.py tests/test_gpu_rollout.py tests/test_gpu_step_host.py -x -q -m gpu 2>&1 | tail -4
for v in 1 0 1 0; do
  export DRL_FAST_SPLIT=$v
  timeout 200 python bench.py --config c4 --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --no-self-check 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('split=$v C4', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],4))"
done
