#!/bin/bash
# quick A/B of warp-kernel builds at the headline shape (no parity run): kernel ms / fraction
for so in "$@"; do
  export DRONERL_B200_LIB=$PWD/$so
  timeout 120 python bench.py --steps 60 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$so G9n4', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],4))"
done
