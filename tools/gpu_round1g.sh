#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_synthesis.py -m gpu -x -q -k "lm or solver" > gpurun_out/g_pytest.log 2>&1; tail -3 gpurun_out/g_pytest.log
QSB_LM_PROFILE=1 timeout 600 python tools/solve_bench.py --method lm > gpurun_out/g_solve_lm2_prof.log 2>&1; grep -E "instances=[0-9]{2,}|case" gpurun_out/g_solve_lm2_prof.log | cut -c1-250
