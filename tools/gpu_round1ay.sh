#!/bin/bash
# LM solver kernel with a warp-uniform warp index (broadcast of lane 0): GPU suite, solver latency / phase profile, synthesis
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/ay_pytest_gpu.log 2>&1; tail -4 gpurun_out/ay_pytest_gpu.log
QSB_LM_PROFILE=1 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/ay_solve_lm2_prof.log 2>&1; grep -E "case" gpurun_out/ay_solve_lm2_prof.log | cut -c1-200
timeout 200 python tools/solve_bench.py --method lm > gpurun_out/ay_solve_lm2.log 2>&1; grep -E "case" gpurun_out/ay_solve_lm2.log | cut -c1-200
timeout 300 python tools/synth_bench.py --cases qft3,toffoli,qft4_leap --reps 2 > gpurun_out/ay_synth_defaults.log 2>&1; cut -c1-220 gpurun_out/ay_synth_defaults.log
timeout 500 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 2 --timeout 200 > gpurun_out/ay_synth_qft5.log 2>&1; cut -c1-260 gpurun_out/ay_synth_qft5.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ay_smoke.log 2>&1; tail -2 gpurun_out/ay_smoke.log
