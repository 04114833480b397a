#!/bin/bash
# C4 (large world) bench line, summarised: ms/step, kernel ms, fraction of HBM, drone-steps/s
python bench.py --grid 64 --drones 32 --radius 5 --stations-factor 8 --skyscrapers-factor 16 --envs-per-gpu ${1:-262144} --steps 20 --warmup 3 --no-cpu-baseline --no-e2e | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('C4', d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['value'])"
