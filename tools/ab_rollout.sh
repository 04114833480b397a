#!/bin/bash
# rollout ms/step (value path) at the headline shape for each library given
for so in "$@"; do
  export DRONERL_B200_LIB=$PWD/$so
  timeout 200 python bench.py --no-cpu-baseline --no-others --no-e2e --no-self-check --steps 99 --warmup 1 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$so', 'ms_per_step(no reset)', round(d['ms_per_step'],4), 'kernel_ms', round(d['roofline']['kernel_ms'],4))"
done
