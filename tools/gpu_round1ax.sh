#!/bin/bash
# ncu of the HS kernel with one CTA per instance (n=4, n=5)
set -x
mkdir -p gpurun_out
for n in 4 5; do
  inst=$([ $n = 4 ] && echo 40000 || echo 10000)
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_hs -c 1 -o /tmp/ax_n$n python tools/prof_case.py --what hs --qubits $n --instances $inst --circuits 100 --reps 2 > gpurun_out/ax_ncu_full_hs_n$n.log 2>&1
  ncu -i /tmp/ax_n$n.ncu-rep --page raw --csv > gpurun_out/ax_raw_hs_n$n.csv 2>/dev/null
  ncu -i /tmp/ax_n$n.ncu-rep --page source --csv > gpurun_out/ax_src_hs_n$n.csv 2>/dev/null
done
