#!/bin/bash
mkdir -p gpurun_out
for i in 1 2; do timeout 600 python tools/solve_bench.py --method lm 2>&1 | cut -c1-200 | sed "s/^/smem$i /"; done
QSB_LM_CHOL=global timeout 600 python tools/solve_bench.py --method lm 2>&1 | cut -c1-200 | sed "s/^/global /"
