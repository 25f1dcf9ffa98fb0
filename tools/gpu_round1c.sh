#!/bin/bash
# third GPU call: DMMA warp kernels -- parity, bench n=4 / n=5, ncu launch list + full capture
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 --deselect tests/test_gpu_synthesis.py::test_qft4_leap > gpurun_out/c_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/c_pytest_gpu.log
for q in 3 4 5; do for w in resid matjac; do
timeout 120 python tools/prof_case.py --what $w --qubits $q --instances 20000 --reps 4 2>&1 | tee -a gpurun_out/c_prof_cases.log
QSB_KERNEL=generic timeout 120 python tools/prof_case.py --what $w --qubits $q --instances 20000 --reps 4 2>&1 | sed 's/^/generic: /' | tee -a gpurun_out/c_prof_cases.log
done; done
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c_bench_n4.json 2> gpurun_out/c_bench_n4.err; echo "bench rc=$?"
cat gpurun_out/c_bench_n4.json; tail -3 gpurun_out/c_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/c_bench_n5.json 2> gpurun_out/c_bench_n5.err; echo "bench5 rc=$?"
cat gpurun_out/c_bench_n5.json; tail -3 gpurun_out/c_bench_n5.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c_launches_resid_n4.csv python tools/prof_case.py --what resid --qubits 4 > gpurun_out/c_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o gpurun_out/c_prof_resid_n4_warp python tools/prof_case.py --what resid --qubits 4 --instances 20000 --reps 2 > gpurun_out/c_ncu_full.log 2>&1
tail -3 gpurun_out/c_ncu_full.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o gpurun_out/c_prof_resid_n5_warp python tools/prof_case.py --what resid --qubits 5 --instances 8000 --reps 2 > gpurun_out/c_ncu_full5.log 2>&1
ls -la gpurun_out
