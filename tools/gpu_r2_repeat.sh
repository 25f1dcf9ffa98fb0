#!/bin/bash
# run-to-run spread of identical LM solves, with the SM clock sampled alongside: tools/gpu_r2_repeat.sh <tag>
tag=$1
mkdir -p gpurun_out
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 100 > gpurun_out/${tag}_clocks.csv &
SMI=$!
for c in "n5 d30 6x4" "n4 d10 4x250" "n4 d8 3x4"; do
  timeout 300 python tools/solve_bench.py --method lm --reps 10 --only "$c" >> gpurun_out/${tag}_repeat.log 2>&1
done
kill $SMI
grep case gpurun_out/${tag}_repeat.log | cut -c1-300
sort gpurun_out/${tag}_clocks.csv | uniq -c | sort -rn | head -8
