#!/bin/bash
# half-size rank-one products + dynamic k layout in the hot kernel: parity, bench n=4/n=5, LM profile, ncu full capture
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s_smoke.log 2>&1; tail -4 gpurun_out/s_smoke.log
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_synthesis.py::test_qft4_leap > gpurun_out/s_pytest_gpu.log 2>&1; tail -5 gpurun_out/s_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/s_bench_n4.json 2> gpurun_out/s_bench_n4.err; cut -c1-200 gpurun_out/s_bench_n4.json; tail -3 gpurun_out/s_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/s_bench_n5.json 2> gpurun_out/s_bench_n5.err; cut -c1-200 gpurun_out/s_bench_n5.json
QSB_LM_PROFILE=1 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/s_solve_lm2_prof.log 2>&1; grep -E "instances=[0-9]{2,}|case" gpurun_out/s_solve_lm2_prof.log | cut -c1-250
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o gpurun_out/s_prof_resid_n4_warp python tools/prof_case.py --what resid --qubits 4 --instances 20000 --reps 2 > gpurun_out/s_ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o gpurun_out/s_prof_resid_n5_warp python tools/prof_case.py --what resid --qubits 5 --instances 4000 --reps 2 > gpurun_out/s_ncu_full5.log 2>&1
ls -la gpurun_out | grep " s_"
