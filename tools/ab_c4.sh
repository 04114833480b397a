#!/bin/bash
# A/B the CTA-mode step kernel at C4: parity of the 64x64 / 12x12 / 32x32 cases, then kernel ms / roofline fraction
for so in "$@"; do
  export DRONERL_B200_LIB=$PWD/$so
  timeout 90 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "64-32-5 or 12-6-4 or 32-16-3 or 12-6-2 or 20-32-2 or 16-4-7 or 24-3-11" 2>&1 | tail -1
  timeout 90 python bench.py --grid 64 --drones 32 --radius 5 --stations-factor 8 --skyscrapers-factor 16 --envs-per-gpu ${ENVS:-131072} --steps 12 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$so', 'kernel_ms', round(d['roofline']['kernel_ms'],3), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],3))"
done
