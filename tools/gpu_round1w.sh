#!/bin/bash
# LM2: split-K SYRK, straight-line diagonal block, column-oriented panel, register back-substitution; then synthesis
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_synthesis.py -m gpu -x -q -k "lm or solver or system" > gpurun_out/w_pytest_lm.log 2>&1; tail -3 gpurun_out/w_pytest_lm.log
QSB_LM_PROFILE=1 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/w_solve_lm2_prof.log 2>&1; grep -E "instances=[0-9]{2,}|case" gpurun_out/w_solve_lm2_prof.log | cut -c1-250
timeout 600 python tools/synth_bench.py --cases qft3,toffoli,fredkin,peres,or,qft3_ring,toffoli_ring,qft4_leap --reps 3 > gpurun_out/w_synth_defaults.log 2>&1; cat gpurun_out/w_synth_defaults.log | cut -c1-250
timeout 700 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 2 --timeout 330 > gpurun_out/w_synth_qft5.log 2>&1; cat gpurun_out/w_synth_qft5.log | cut -c1-300
