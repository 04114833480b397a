#!/bin/bash
# session GPU pass: new trainer tests, C5 bench, write ceiling, C4 bench + ncu capture
mkdir -p gpurun_out
python -m pytest tests/test_gpu_trainer.py -x -q -m gpu > gpurun_out/pytest_trainer.log 2>&1; echo "trainer tests rc=$?"
tail -15 gpurun_out/pytest_trainer.log
python tools/bench_c5.py --steps 1000 --warmup 100 > gpurun_out/c5_n1.json 2> gpurun_out/c5_n1.err; tail -2 gpurun_out/c5_n1.json gpurun_out/c5_n1.err
python tools/bench_c5.py --steps 300 --warmup 50 --no-graph > gpurun_out/c5_n1_nograph.json 2>> gpurun_out/c5_n1.err; tail -1 gpurun_out/c5_n1_nograph.json
python tools/write_ceiling.py > gpurun_out/write_ceiling.txt 2>&1; cat gpurun_out/write_ceiling.txt
bash tools/c4.sh > gpurun_out/c4_now.txt 2>&1; cat gpurun_out/c4_now.txt
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_step_cta -s 3 -c 1 -o gpurun_out/step_cta_r1b -f \
  python bench.py --grid 64 --drones 32 --radius 5 --stations-factor 8 --skyscrapers-factor 16 --envs-per-gpu 32768 --steps 3 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/ncu_c4b.log 2>&1
echo "ncu rc=$?"
