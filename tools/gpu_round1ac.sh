#!/bin/bash
# validation of the state with the StateprepObjective: smoke, full GPU suite (incl. tests/test_stateprep.py), benches n=4 / n=5, reference arm
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ac_smoke.log 2>&1; tail -4 gpurun_out/ac_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/ac_pytest_gpu.log 2>&1; tail -14 gpurun_out/ac_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/ac_bench_n4.json 2> gpurun_out/ac_bench_n4.err; cut -c1-200 gpurun_out/ac_bench_n4.json; tail -3 gpurun_out/ac_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/ac_bench_n5.json 2> gpurun_out/ac_bench_n5.err; cut -c1-200 gpurun_out/ac_bench_n5.json; tail -3 gpurun_out/ac_bench_n5.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/ac_bench_reference.json 2>&1; cut -c1-200 gpurun_out/ac_bench_reference.json
ls gpurun_out | grep "^ac_"
