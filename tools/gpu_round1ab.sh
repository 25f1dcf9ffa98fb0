#!/bin/bash
# ncu full captures of the solver kernel (frontier-shaped n=5 batch) and of the LM-system kernel; pages exported on the box
set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qsb_k_lm2 --launch-skip 1 -c 1 -o /tmp/ab_lm2 python tools/solve_bench.py --method lm --only "n5 d30 6x4" > gpurun_out/ab_ncu_lm2.log 2>&1; tail -3 gpurun_out/ab_ncu_lm2.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_lmsys -c 1 -o /tmp/ab_lmsys python tools/prof_case.py --what lmsys --qubits 4 --instances 5000 --reps 2 > gpurun_out/ab_ncu_lmsys.log 2>&1; tail -3 gpurun_out/ab_ncu_lmsys.log
for k in lm2 lmsys; do
  ncu -i /tmp/ab_$k.ncu-rep --page raw --csv > gpurun_out/ab_raw_$k.csv 2>/dev/null
  ncu -i /tmp/ab_$k.ncu-rep --page source --csv > gpurun_out/ab_src_$k.csv 2>/dev/null
done
ls -la gpurun_out | grep " ab_"
