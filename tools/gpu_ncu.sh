#!/bin/bash
# ncu evidence: launch list of the bench command + --set full captures of the dominant kernels (bench ran clean first).
# The .ncu-rep files are summarised ON THE BOX (tools/ncu_breakdown.py + raw/source csv) and deleted: together they
# exceed the 64 MiB that gpurun copies back.
mkdir -p gpurun_out
TAG=${1:-r02}
WHAT=${2:-all}
cap() {  # name nenv kernel-regex skip count cmd...
  name=$1; nenv=$2; rx=$3; skip=$4; cnt=$5; shift 5
  timeout 400 ncu --set full --import-source on --clock-control none -k regex:$rx -s $skip -c $cnt -o gpurun_out/${TAG}_${name} -f "$@" > gpurun_out/ncu_${name}.log 2>&1; echo "ncu $name rc=$?"
  python tools/ncu_breakdown.py gpurun_out/${TAG}_${name}.ncu-rep $nenv > gpurun_out/${TAG}_${name}_ncu_full.txt 2>&1
  ncu -i gpurun_out/${TAG}_${name}.ncu-rep --page raw --csv 2>/dev/null | gzip > gpurun_out/${TAG}_${name}_raw.csv.gz
  python tools/ncu_lines.py gpurun_out/${TAG}_${name}.ncu-rep '' 90 > gpurun_out/${TAG}_${name}_lines.txt 2>&1
  rm -f gpurun_out/${TAG}_${name}.ncu-rep
}
if [ $WHAT = all ]; then
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-others > gpurun_out/${TAG}_bench_short.json 2> gpurun_out/${TAG}_bench_short.err; echo "bench rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-others --no-self-check > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
fi
B="python bench.py --no-cpu-baseline --no-e2e --no-others --no-self-check"
if [ $WHAT = all -o $WHAT = warp ]; then cap step_warp 1048576 k_step_warp 8 1 $B --steps 8 --warmup 5; fi
if [ $WHAT = all ]; then
cap obs_warp 1048576 k_obs_warp 2 1 python tools/time_obs.py
cap reset_warp 1048576 k_reset_warp 1 1 $B --steps 101 --warmup 5
cap derive 1048576 k_derive 3 1 $B --steps 8 --warmup 5
cap step_g8n3 1048576 k_step_warp 8 1 $B --config g8n3 --steps 8 --warmup 5
fi
# C4 in steady state: the 45th fast-kernel launch after the reset
if [ $WHAT = all -o $WHAT = c4 ]; then
cap step_c4 262144 k_step_fast 45 1 $B --config c4 --steps 30 --warmup 20
fi
ls -la gpurun_out/ | tail -24
