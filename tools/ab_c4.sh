#!/bin/bash
# A/B the C4 step kernels for each library given: parity on the CTA-mode shapes, then kernel ms at C4
for so in "$@"; do
  export DRONERL_B200_LIB=$PWD/$so
  timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "rollout_matches_oracle and (12- or 64- or 20-32 or 16-4-7 or 32-16 or 24-3 or 90-8)" 2>&1 | tail -1
  timeout 200 python bench.py --config c4 --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --no-self-check 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$so C4', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],4))"
done
