#!/bin/bash
# wall-time repeatability of the solver benchmark (no profiling), clocks sampled around it
mkdir -p gpurun_out
q="--query-gpu=clocks.sm,clocks.max.sm,clocks_throttle_reasons.active,power.draw,temperature.gpu --format=csv,noheader"
for rep in 1 2 3; do
  nvidia-smi $q
  timeout 200 python tools/solve_bench.py --method lm 2>&1 | grep case | cut -c1-200
done > gpurun_out/y_solve_repeat.log 2>&1
nvidia-smi $q >> gpurun_out/y_solve_repeat.log
cat gpurun_out/y_solve_repeat.log
