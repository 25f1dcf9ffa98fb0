#!/bin/bash
# ncu of the n=5 hot kernel: 4-warp CTAs (v1) vs one-warp CTAs (exp)
set -x
mkdir -p gpurun_out
for lib in v1 exp; do
QSB_LIB_PATH=qsearch_b200/libqsb200_$lib.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o /tmp/ak_$lib python tools/prof_case.py --what resid --qubits 5 --instances 4000 --reps 2 > gpurun_out/ak_ncu_$lib.log 2>&1
ncu -i /tmp/ak_$lib.ncu-rep --page raw --csv > gpurun_out/ak_raw_$lib.csv 2>/dev/null
ncu -i /tmp/ak_$lib.ncu-rep --page source --csv > gpurun_out/ak_src_$lib.csv 2>/dev/null
done
