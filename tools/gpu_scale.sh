#!/bin/bash
# bench line (weak + strong block) at N GPUs: tools/gpu_scale.sh N [tag]
mkdir -p gpurun_out
NP=${1:-2}; TAG=${2:-r02}
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NP > gpurun_out/${TAG}_bench_n$NP.json 2> gpurun_out/${TAG}_bench_n$NP.err; echo "bench n=$NP rc=$?"
python - <<PY
import json
line=[l for l in open("gpurun_out/${TAG}_bench_n$NP.json") if l.startswith("{")][-1]   # (NCCL prints its version banner on stdout first)
open("gpurun_out/${TAG}_bench_n$NP.json","w").write(line)
d=json.loads(line)
s=d.get("strong") or {}
print("N=$NP weak value %.4g ms/step %.4f e2e %.4g | strong value %.4g ms/step %.4f step/kernel %.3f e2e %.4g" % (d["value"], d["ms_per_step"], d["e2e"]["value"], s.get("value",0), s.get("ms_per_step",0), s.get("step_over_kernel",0), (s.get("e2e") or {}).get("value",0)))
PY
