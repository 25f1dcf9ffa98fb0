#!/bin/bash
# LM damping schedule experiments (qsb_k_lm3): attempts per Jacobian evaluation, iterations, wall.  tools/gpu_r2_mu.sh <tag>
tag=$1
mkdir -p gpurun_out
run() {
  echo "== $1"
  for c in "n4 d10 4x250" "n4 d8 3x4" "n5 d12 6x4" "n3 d6 2x4"; do
    env $1 QSB_LM_PROFILE=1 timeout 100 python tools/solve_bench.py --method lm --only "$c" 2>&1 | grep -h "instances=[0-9]*[0-9][0-9] \|instances=8 \|case" | cut -c1-110,230-330
  done
}
{
run "QSB_X=0"
run "QSB_LM_MU_DOWN=0.5"
run "QSB_LM_MU_DOWN=0.7"
run "QSB_LM_MU_STREAK=1"
run "QSB_LM_MU_STREAK=2"
run "QSB_LM_MU_UP=3"
run "QSB_LM_MU_DOWN=0.5 QSB_LM_MU_STREAK=1"
} > gpurun_out/${tag}_mu_schedule.log 2>&1
cat gpurun_out/${tag}_mu_schedule.log
