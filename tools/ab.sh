#!/bin/bash
# A/B the step kernel: run bench.py with each library in variants/*.so (kernel_ms, frac)
for so in variants/*.so; do
  DRONERL_B200_LIB=$PWD/$so python bench.py --steps 60 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$so', round(d['roofline']['kernel_ms'],4), round(d['roofline']['frac'],4), round(d['ms_per_step'],4))"
done
