#!/bin/bash
# hot kernel: ops staged in shared memory, normalised rank-one factor, one dispatch per op
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_synthesis.py::test_qft4_leap > gpurun_out/u_pytest_gpu.log 2>&1; tail -5 gpurun_out/u_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/u_bench_n4.json 2> gpurun_out/u_bench_n4.err; cut -c1-200 gpurun_out/u_bench_n4.json; tail -3 gpurun_out/u_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 --no-cpu > gpurun_out/u_bench_n5.json 2> gpurun_out/u_bench_n5.err; cut -c1-200 gpurun_out/u_bench_n5.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o /tmp/u_n4 python tools/prof_case.py --what resid --qubits 4 --instances 20000 --reps 2 > gpurun_out/u_ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o /tmp/u_n5 python tools/prof_case.py --what resid --qubits 5 --instances 4000 --reps 2 > gpurun_out/u_ncu_full5.log 2>&1
for n in 4 5; do
  ncu -i /tmp/u_n$n.ncu-rep --page raw --csv > gpurun_out/u_raw_n$n.csv 2>/dev/null
  ncu -i /tmp/u_n$n.ncu-rep --page source --csv > gpurun_out/u_src_n$n.csv 2>/dev/null
done
ls -la gpurun_out | grep " u_"
