#!/bin/bash
# LM2 with vectorised SYRK loads: parity of the solver tests + phase profile
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lm or solver or system" > gpurun_out/v_pytest_lm.log 2>&1; tail -3 gpurun_out/v_pytest_lm.log
QSB_LM_PROFILE=1 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/v_solve_lm2_prof.log 2>&1; grep -E "instances=[0-9]{2,}|case" gpurun_out/v_solve_lm2_prof.log | cut -c1-250
