#!/bin/bash
# kernel ms at C4 for each library given (no parity run)
for so in "$@"; do
  export DRONERL_B200_LIB=$PWD/$so
  timeout 200 python bench.py --config c4 --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --no-self-check 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$so C4', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],4))"
done
