#!/bin/bash
# LM solver A/B on frontier-shaped batches: tools/gpu_r2_solve.sh <tag>   (general kernel qsb_k_lm2 vs body-frame qsb_k_lm3)
tag=$1
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_gpus.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/${tag}_pytest_gpu.log 2>&1; tail -12 gpurun_out/${tag}_pytest_gpu.log
for k in lm2 lm3; do
  QSB_LM_KERNEL=$k QSB_LM_PROFILE=1 timeout 300 python tools/solve_bench.py --method lm > gpurun_out/${tag}_solve_${k}_prof.log 2>&1
  QSB_LM_KERNEL=$k timeout 300 python tools/solve_bench.py --method lm --reps 3 > gpurun_out/${tag}_solve_${k}.log 2>&1
  grep case gpurun_out/${tag}_solve_${k}.log | cut -c1-260
done
grep profile gpurun_out/${tag}_solve_lm3_prof.log | cut -c1-300
timeout 300 python tools/solve_bench.py --method lbfgs --reps 2 > gpurun_out/${tag}_solve_lbfgs.log 2>&1; grep case gpurun_out/${tag}_solve_lbfgs.log | cut -c1-200
