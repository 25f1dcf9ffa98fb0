#!/bin/bash
# HS warp kernel + L-BFGS v2: full GPU suite, bench, solver latency
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/n_pytest_gpu.log 2>&1; tail -9 gpurun_out/n_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/n_bench_n4.json 2> gpurun_out/n_bench_n4.err; cut -c1-200 gpurun_out/n_bench_n4.json; tail -3 gpurun_out/n_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/n_bench_n5.json 2> gpurun_out/n_bench_n5.err; cut -c1-200 gpurun_out/n_bench_n5.json
timeout 300 python tools/solve_bench.py --method lbfgs > gpurun_out/n_solve_lbfgs2.log 2>&1; cut -c1-260 gpurun_out/n_solve_lbfgs2.log
timeout 600 python tools/synth_bench.py --cases qft3,toffoli,qft4_leap --reps 2 --solver lbfgs --starts 4 > gpurun_out/n_synth_lbfgs2.log 2>&1; cut -c1-260 gpurun_out/n_synth_lbfgs2.log
for q in 4 5; do timeout 120 python tools/prof_case.py --what hs --qubits $q --instances 100000 --reps 4 2>&1 | tee -a gpurun_out/n_prof_hs.log; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_hs -c 1 -o gpurun_out/n_prof_hs_n4_warp python tools/prof_case.py --what hs --qubits 4 --instances 100000 --reps 2 > gpurun_out/n_ncu_hs.log 2>&1
