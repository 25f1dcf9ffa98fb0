#!/bin/bash
# second GPU call: full GPU test suite (incl. on-device solvers + synthesis) and end-to-end synthesis timings
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/b_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/b_pytest_gpu.log
timeout 900 python tools/synth_bench.py --cases qft3,toffoli,toffoli_ring,qft4_leap --reps 2 --starts 4 > gpurun_out/b_synth.log 2>&1; echo "synth rc=$?"
cat gpurun_out/b_synth.log
timeout 900 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 1 --starts 4 --timeout 600 > gpurun_out/b_synth_qft5.log 2>&1; echo "synth5 rc=$?"
cat gpurun_out/b_synth_qft5.log
