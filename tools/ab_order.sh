#!/bin/bash
# A/B of the step kernels' env order: tickets (default) vs static grid-stride order (DRL_STATIC_ORDER=1)
mkdir -p gpurun_out
line() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],4), 'ms_per_step', round(d['ms_per_step'],4), 'value', '%.4g' % d['value'], 'e2e', '%.4g' % (d.get('e2e') or {}).get('value', 0))"; }
C4="--grid 64 --drones 32 --radius 5 --stations-factor 8 --skyscrapers-factor 16 --envs-per-gpu 262144 --steps 20 --warmup 3 --no-cpu-baseline"
for so in 0 1; do
  export DRL_STATIC_ORDER=$so
  timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline 2>&1 | line "G9n4 static=$so"
  timeout 200 python bench.py --grid 8 --drones 3 --steps 100 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | line "G8n3 static=$so"
  timeout 200 python bench.py $C4 2>&1 | line "C4 static=$so"
done
