#!/bin/bash
# straggler / sibling stop rules on the frontier-shaped batches: tools/gpu_r2_strag.sh <tag>
tag=$1
mkdir -p gpurun_out
for c in 0 2 4; do
  timeout 300 python tools/solve_bench.py --method lm --reps 3 --straggler $c > gpurun_out/${tag}_solve_lm3_strag$c.log 2>&1
  grep case gpurun_out/${tag}_solve_lm3_strag$c.log | cut -c1-330
done
