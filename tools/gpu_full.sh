#!/bin/bash
# full GPU pass: whole -m gpu suite, smoke, headline bench (+ reference arm), launch list, C4 + G8n3 lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_full.log
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 300 python bench.py > gpurun_out/bench_main.json 2> gpurun_out/bench_main.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_main.json
timeout 200 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 300 python bench.py --grid 64 --drones 32 --radius 5 --stations-factor 8 --skyscrapers-factor 16 --envs-per-gpu 262144 --steps 20 --warmup 3 --cpu-sample-envs 2048 --cpu-seconds 8 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "c4 rc=$?"
timeout 300 python bench.py --grid 8 --drones 3 --no-cpu-baseline > gpurun_out/bench_g8n3.json 2> gpurun_out/bench_g8n3.err; echo "g8n3 rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
# ncu --set full captures of the dominant kernels at the bench shapes (DRAM traffic per launch -> profiles/step_kernel_traffic.json)
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_step_warp -s 5 -c 1 -o gpurun_out/step_warp_full -f python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_warp_full.log 2>&1; echo "ncu warp rc=$?"
timeout 400 ncu --set full --import-source on --clock-control none -k regex:k_step_ -s 6 -c 2 -o gpurun_out/step_c4_full -f python bench.py --grid 64 --drones 32 --radius 5 --stations-factor 8 --skyscrapers-factor 16 --envs-per-gpu 262144 --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_c4_full.log 2>&1; echo "ncu c4 rc=$?"
timeout 200 python tools/bench_c5.py --steps 1000 --warmup 100 > gpurun_out/c5_n1.json 2> gpurun_out/c5_n1.err; echo "c5 rc=$?"; tail -c 400 gpurun_out/c5_n1.json
