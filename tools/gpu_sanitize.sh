#!/bin/bash
# compute-sanitizer over every kernel family at three shapes (warp mode, CTA mode fast+redo, the C4 large world)
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
run() {  # tool shape...
  tool=$1; shift; tag=$1; shift
  timeout 900 $CS --tool $tool --print-limit 20 python tools/sanitize_run.py "$@" > gpurun_out/sanitize_${tool}_${tag}.log 2>&1
  echo "$tool $tag rc=$? :: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|sanitize_run ok' gpurun_out/sanitize_${tool}_${tag}.log | tr '\n' ' ')"
}
for tool in memcheck racecheck synccheck; do
  run $tool g9n4 9 4 3 256
  run $tool g8n3 8 3 3 256
  run $tool g12n6 12 6 2 64
  run $tool g12n6pf8 12 6 2 64 8 2 2 3
  run $tool c4 64 32 5 8 3 2 8 16
done
