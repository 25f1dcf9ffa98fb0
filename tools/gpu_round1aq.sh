#!/bin/bash
# final-state validation: CTA-per-instance kernels with SM-local instance order; smoke, GPU suite, benches, reference arm,
# launch list, ncu of the hot kernel on the bench's 100-circuit shape, HS A/B against the previous build
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/aq_smoke.log 2>&1; tail -4 gpurun_out/aq_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/aq_pytest_gpu.log 2>&1; tail -9 gpurun_out/aq_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/aq_bench_n4.json 2> gpurun_out/aq_bench_n4.err; cut -c1-200 gpurun_out/aq_bench_n4.json; tail -3 gpurun_out/aq_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/aq_bench_n5.json 2> gpurun_out/aq_bench_n5.err; cut -c1-200 gpurun_out/aq_bench_n5.json; tail -3 gpurun_out/aq_bench_n5.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/aq_bench_reference.json 2>&1; cut -c1-200 gpurun_out/aq_bench_reference.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/aq_launches_bench_n4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/aq_ncu_list.log 2>&1
for n in 4 5; do
  inst=$([ $n = 4 ] && echo 20000 || echo 4000)
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o /tmp/aq_n$n python tools/prof_case.py --what resid --qubits $n --instances $inst --circuits 100 --reps 2 > gpurun_out/aq_ncu_full_n$n.log 2>&1
  ncu -i /tmp/aq_n$n.ncu-rep --page raw --csv > gpurun_out/aq_raw_n$n.csv 2>/dev/null
  ncu -i /tmp/aq_n$n.ncu-rep --page source --csv > gpurun_out/aq_src_n$n.csv 2>/dev/null
done
bash tools/gpu_ab_hs.sh aq libqsb200_v2.so libqsb200.so
bash tools/gpu_ab2.sh aq libqsb200_v1.so libqsb200_v2.so libqsb200.so
