#!/bin/bash
mkdir -p gpurun_out
QSB_LM_PROFILE=1 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/q_solve_lm2_prof.log 2>&1; grep -E "instances=[0-9]{2,}" gpurun_out/q_solve_lm2_prof.log | cut -c1-300
