#!/bin/bash
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_trainer.py -x -q -m gpu 2>&1 | tail -3
for f in 0 1; do
DRL_NO_FUSE=$f timeout 120 python bench.py --steps 100 --warmup 10 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('NO_FUSE=$f G9n4 value %.4e ms/step %.4f kernel_ms %.4f e2e %.4e launches %d' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value'], d['gpu_launches']))"
done
DRL_NO_FUSE=0 timeout 200 python bench.py --grid 64 --drones 32 --radius 5 --stations-factor 8 --skyscrapers-factor 16 --envs-per-gpu 262144 --steps 20 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('C4 value %.4e ms/step %.4f kernel_ms %.4f frac %.4f e2e %.4e launches %d' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['e2e']['value'], d['gpu_launches']))"
