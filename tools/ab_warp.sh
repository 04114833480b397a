#!/bin/bash
# A/B the warp-mode step kernel at the headline shape (and the dataclass default) for each library given
for so in "$@"; do
  export DRONERL_B200_LIB=$PWD/$so
  timeout 200 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_gpu_reference_cases.py -x -q -m gpu 2>&1 | tail -1
  timeout 120 python bench.py --steps 60 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$so G9n4', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],4))"
  timeout 120 python bench.py --grid 8 --drones 3 --steps 60 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$so G8n3', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],4))"
done
