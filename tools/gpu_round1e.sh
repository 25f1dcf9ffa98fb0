#!/bin/bash
# fifth GPU call: tensor-core LM (qsb_k_lm2) -- parity tests, solver latency, end-to-end synthesis
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 --deselect tests/test_gpu_synthesis.py::test_qft4_leap > gpurun_out/e_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/e_pytest_gpu.log
timeout 600 python tools/solve_bench.py --method lm > gpurun_out/e_solve_lm2.log 2>&1; cat gpurun_out/e_solve_lm2.log
QSB_KERNEL=generic timeout 900 python tools/solve_bench.py --method lm > gpurun_out/e_solve_lm_generic.log 2>&1; cat gpurun_out/e_solve_lm_generic.log
timeout 300 python tools/solve_bench.py --method lbfgs > gpurun_out/e_solve_lbfgs.log 2>&1; cat gpurun_out/e_solve_lbfgs.log
timeout 900 python tools/synth_bench.py --cases qft3,toffoli,qft4_leap --reps 2 --starts 4 > gpurun_out/e_synth_lm2.log 2>&1; cat gpurun_out/e_synth_lm2.log
timeout 500 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 1 --starts 4 --timeout 400 > gpurun_out/e_synth_qft5_lm2.log 2>&1; cat gpurun_out/e_synth_qft5_lm2.log
