#!/bin/bash
# round-2 first GPU pass: new tests, bench line, e2e pipelines A/B at N=1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_rollout.py tests/test_gpu_step_host.py tests/test_gpu_state_export.py tests/test_gpu_evaluator.py -x -q -m gpu > gpurun_out/pytest_new.log 2>&1; echo "pytest new rc=$?"; tail -15 gpurun_out/pytest_new.log
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_rollout.py --deselect tests/test_gpu_step_host.py --deselect tests/test_gpu_state_export.py --deselect tests/test_gpu_evaluator.py > gpurun_out/pytest_old.log 2>&1; echo "pytest old rc=$?"; tail -3 gpurun_out/pytest_old.log
timeout 400 python bench.py > gpurun_out/bench_main.json 2> gpurun_out/bench_main.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_main.err; python -c "
import json; d=json.load(open('gpurun_out/bench_main.json'))
print('value %.4g ms/step %.4f kernel_ms %.4f frac %.4f e2e %.4g (%s enq %.3f sync %.3f)' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['e2e']['value'], d['e2e']['pipeline'], d['e2e']['host_enqueue_ms_per_call'], d['e2e']['host_sync_wait_ms_per_call']))
for k,v in d.get('other_configs',{}).items(): print(k, '%.4g' % v['value'], 'kernel_ms %.4f frac %.4f' % (v['roofline']['kernel_ms'], v['roofline']['frac']), 'e2e %.4g' % v['e2e']['value'])
"
DRL_SPLIT_DERIVE=1 timeout 200 python bench.py --no-cpu-baseline --no-e2e --no-others > gpurun_out/bench_splitderive.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/bench_splitderive.json')); print('split-derive: ms/step %.4f kernel %.4f' % (d['ms_per_step'], d['roofline']['kernel_ms']))"
timeout 200 python bench.py --scaling strong --envs-per-gpu 131072 --no-cpu-baseline --no-others --steps 400 --warmup 20 > gpurun_out/bench_131k.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/bench_131k.json')); print('131k: ms/step %.4f kernel %.4f ratio %.3f e2e %.4g' % (d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['step_over_kernel'], d['e2e']['value']))"
timeout 300 python tools/e2e_limiter.py --steps 40 > gpurun_out/e2e_limiter_n1.json 2> gpurun_out/e2e_limiter_n1.err; echo "limiter rc=$?"; cat gpurun_out/e2e_limiter_n1.json; tail -3 gpurun_out/e2e_limiter_n1.err
timeout 200 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -c 900 gpurun_out/bench_ref.json
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1; nproc; cat /sys/bus/pci/devices/*/numa_node 2>/dev/null | sort | uniq -c | head
