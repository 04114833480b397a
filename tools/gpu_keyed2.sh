#!/bin/bash
for f in 0 1; do
DRL_NO_FUSE=$f timeout 120 python bench.py --steps 100 --warmup 10 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('NO_FUSE=$f G9n4 value %.4e ms/step %.4f kernel_ms %.4f e2e %.4e launches %d' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value'], d['gpu_launches']))"
done
