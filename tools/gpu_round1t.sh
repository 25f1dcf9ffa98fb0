#!/bin/bash
# ncu full capture of the half-product hot kernel (n=4, n=5); pages exported on the box, only the n=4 report comes back
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o /tmp/t_n4 python tools/prof_case.py --what resid --qubits 4 --instances 20000 --reps 2 > gpurun_out/t_ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o /tmp/t_n5 python tools/prof_case.py --what resid --qubits 5 --instances 4000 --reps 2 > gpurun_out/t_ncu_full5.log 2>&1
for n in 4 5; do
  ncu -i /tmp/t_n$n.ncu-rep --page raw --csv > gpurun_out/t_raw_n$n.csv 2>/dev/null
  ncu -i /tmp/t_n$n.ncu-rep --page source --csv > gpurun_out/t_src_n$n.csv 2>/dev/null
done
cp /tmp/t_n4.ncu-rep gpurun_out/t_prof_resid_n4_warp.ncu-rep
ls -la gpurun_out
