#!/bin/bash
# 8-GPU pass: what limits drl_step_host with 8 ranks on one host, then the bench line (weak + strong block)
mkdir -p gpurun_out
NP=${1:-8}
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29511 tools/e2e_limiter.py --steps 40 > gpurun_out/e2e_limiter_n$NP.json 2> gpurun_out/e2e_limiter_n$NP.err; echo "limiter rc=$?"; cat gpurun_out/e2e_limiter_n$NP.json; tail -3 gpurun_out/e2e_limiter_n$NP.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NP > gpurun_out/bench_n$NP.json 2> gpurun_out/bench_n$NP.err; echo "bench rc=$?"; cat gpurun_out/bench_n$NP.json; tail -3 gpurun_out/bench_n$NP.err
nvidia-smi topo -m > gpurun_out/topo_n$NP.txt 2>&1; nproc
