#!/bin/bash
# L-BFGS solver kernel with one warp per column block (uniform control flow): GPU suite, solver latency, stateprep synthesis
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/aw_pytest_gpu.log 2>&1; tail -4 gpurun_out/aw_pytest_gpu.log
timeout 200 python tools/solve_bench.py --method lbfgs > gpurun_out/aw_solve_lbfgs2.log 2>&1; cut -c1-250 gpurun_out/aw_solve_lbfgs2.log
timeout 300 python tools/synth_bench.py --cases qft3,toffoli --solver lbfgs --reps 2 > gpurun_out/aw_synth_lbfgs.log 2>&1; cut -c1-220 gpurun_out/aw_synth_lbfgs.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/aw_smoke.log 2>&1; tail -2 gpurun_out/aw_smoke.log
timeout 600 python bench.py --no-cpu > gpurun_out/aw_bench_n4.json 2> gpurun_out/aw_bench_n4.err; cut -c1-160 gpurun_out/aw_bench_n4.json
