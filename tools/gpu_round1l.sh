#!/bin/bash
# gate table + DMMA reorder: parity, bench n=4/n=5, LM profile, ncu captures (launch list + full) of the hot kernel
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_synthesis.py::test_qft4_leap > gpurun_out/l_pytest_gpu.log 2>&1; tail -3 gpurun_out/l_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/l_bench_n4.json 2> gpurun_out/l_bench_n4.err; cat gpurun_out/l_bench_n4.json | cut -c1-400; tail -3 gpurun_out/l_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/l_bench_n5.json 2> gpurun_out/l_bench_n5.err; cat gpurun_out/l_bench_n5.json | cut -c1-400
QSB_LM_PROFILE=1 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/l_solve_lm2_prof.log 2>&1; grep -E "instances=[0-9]{2,}|case" gpurun_out/l_solve_lm2_prof.log | cut -c1-250
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/l_launches_bench_n4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/l_ncu_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o gpurun_out/l_prof_resid_n4_warp python tools/prof_case.py --what resid --qubits 4 --instances 20000 --reps 2 > gpurun_out/l_ncu_full.log 2>&1
ls -la gpurun_out | grep " l_"
