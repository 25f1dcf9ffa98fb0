#!/bin/bash
# first GPU call: smoke, parity tests, bench (n=4, n=5), ncu launch list + one full capture of the residual-Jacobian kernel
set -x
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
(lscpu | head -20; free -g) > gpurun_out/host.txt 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
tail -5 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err; echo "bench rc=$?"
cat gpurun_out/bench_n4.json; tail -3 gpurun_out/bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/bench_n5.json 2> gpurun_out/bench_n5.err; echo "bench5 rc=$?"
cat gpurun_out/bench_n5.json; tail -3 gpurun_out/bench_n5.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>&1; cat gpurun_out/bench_ref.json
for w in resid hs; do
timeout 120 python tools/prof_case.py --what $w --qubits 4 > gpurun_out/prof_case_$w.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$w.csv python tools/prof_case.py --what $w --qubits 4 > gpurun_out/ncu_list_$w.log 2>&1
cat gpurun_out/prof_case_$w.log
done
timeout 120 python tools/prof_case.py --what resid --qubits 4 --reps 2 > gpurun_out/prof_case2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qsb_k_jac -c 1 -o gpurun_out/prof_resid_n4_r1a python tools/prof_case.py --what resid --qubits 4 --reps 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out
