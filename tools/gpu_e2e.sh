#!/bin/bash
for ch in 4 8 12 15; do
DRL_HOST_CHUNKS=$ch timeout 120 python bench.py --steps 60 --warmup 10 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('chunks=$ch value %.4e e2e %.4e' % (d['value'], d['e2e']['value']))"
done
