#!/bin/bash
# A/B warp-kernel builds: parity suite first, then kernel ms / rollout / e2e at the headline shape and G8n3
line() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],4), 'value', '%.4g' % d['value'], 'e2e', '%.4g' % (d.get('e2e') or {}).get('value', 0))"; }
for so in "$@"; do
  export DRONERL_B200_LIB=$PWD/$so
  timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_gpu_reference_cases.py -x -q -m gpu 2>&1 | tail -1
  timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline 2>&1 | line "$so G9n4"
  timeout 200 python bench.py --grid 8 --drones 3 --steps 100 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | line "$so G8n3"
done
