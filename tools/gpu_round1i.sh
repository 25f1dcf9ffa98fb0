#!/bin/bash
# cluster-split tensor-core LM: parity tests, latency profile, end-to-end synthesis
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_synthesis.py -m gpu -x -q -k "lm or solver" > gpurun_out/i_pytest.log 2>&1; tail -3 gpurun_out/i_pytest.log
QSB_LM_PROFILE=1 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/i_solve_lm2_prof.log 2>&1; grep -E "instances=[0-9]{2,}|case" gpurun_out/i_solve_lm2_prof.log | cut -c1-250
timeout 600 python tools/synth_bench.py --cases qft3,toffoli,qft4_leap --reps 2 --starts 4 > gpurun_out/i_synth_lm2.log 2>&1; cat gpurun_out/i_synth_lm2.log
timeout 500 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 1 --starts 4 --timeout 400 > gpurun_out/i_synth_qft5_lm2.log 2>&1; cat gpurun_out/i_synth_qft5_lm2.log
