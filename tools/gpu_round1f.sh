#!/bin/bash
set -x
mkdir -p gpurun_out
QSB_LM_PROFILE=1 timeout 600 python tools/solve_bench.py --method lm > gpurun_out/f_solve_lm2_prof.log 2>&1; cat gpurun_out/f_solve_lm2_prof.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/f_bench_n4.json 2> gpurun_out/f_bench_n4.err; cat gpurun_out/f_bench_n4.json; tail -3 gpurun_out/f_bench_n4.err
