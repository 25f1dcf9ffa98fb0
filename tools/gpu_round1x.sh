#!/bin/bash
# validation of the state after the half-product hot kernel and the LM solver work: smoke, full GPU suite, benches, LM profile, launch list
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/x_smoke.log 2>&1; tail -4 gpurun_out/x_smoke.log
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/x_pytest_gpu.log 2>&1; tail -9 gpurun_out/x_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/x_bench_n4.json 2> gpurun_out/x_bench_n4.err; cut -c1-200 gpurun_out/x_bench_n4.json; tail -3 gpurun_out/x_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/x_bench_n5.json 2> gpurun_out/x_bench_n5.err; cut -c1-200 gpurun_out/x_bench_n5.json; tail -3 gpurun_out/x_bench_n5.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/x_bench_reference.json 2>&1; cut -c1-200 gpurun_out/x_bench_reference.json
QSB_LM_PROFILE=1 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/x_solve_lm2_prof.log 2>&1; grep -E "instances=[0-9]{2,}|case" gpurun_out/x_solve_lm2_prof.log | cut -c1-250
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/x_launches_bench_n4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/x_ncu_list.log 2>&1
ls gpurun_out | grep "^x_"
