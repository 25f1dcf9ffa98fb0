#!/bin/bash
# fourth GPU call: L-BFGS as the search solver (cheap HS gradient) vs LM, end-to-end
set -x
mkdir -p gpurun_out
timeout 600 python tools/synth_bench.py --cases qft3,toffoli,qft4_leap --reps 2 --starts 4 --solver lbfgs > gpurun_out/d_synth_lbfgs.log 2>&1; echo "rc=$?"
cat gpurun_out/d_synth_lbfgs.log
timeout 500 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 1 --starts 8 --solver lbfgs --timeout 400 > gpurun_out/d_synth_qft5_lbfgs.log 2>&1; echo "rc=$?"
cat gpurun_out/d_synth_qft5_lbfgs.log
