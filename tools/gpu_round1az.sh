#!/bin/bash
# final state of the round (after the solver-kernel changes): smoke, GPU suite, bench lines n=4 / n=5, reference arm, launch list
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/az_smoke.log 2>&1; tail -4 gpurun_out/az_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/az_pytest_gpu.log 2>&1; tail -9 gpurun_out/az_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/az_bench_n4.json 2> gpurun_out/az_bench_n4.err; cut -c1-200 gpurun_out/az_bench_n4.json; tail -3 gpurun_out/az_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/az_bench_n5.json 2> gpurun_out/az_bench_n5.err; cut -c1-200 gpurun_out/az_bench_n5.json; tail -3 gpurun_out/az_bench_n5.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/az_bench_reference.json 2>&1; cut -c1-200 gpurun_out/az_bench_reference.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/az_launches_bench_n4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/az_ncu_list.log 2>&1
QSB_LM_PROFILE=0 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/az_solve_lm2.log 2>&1; grep case gpurun_out/az_solve_lm2.log | cut -c1-170
