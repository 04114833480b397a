#!/bin/bash
# multi-GPU pass: headline bench and C5 trainer at N GPUs (N = first arg)
N=${1:-8}
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 200 --warmup 20 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
tail -c 1500 gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/bench_c5.py --steps 1000 --warmup 100 > gpurun_out/c5_n$N.json 2> gpurun_out/c5_n$N.err
tail -c 800 gpurun_out/c5_n$N.json; tail -3 gpurun_out/c5_n$N.err
