#!/bin/bash
set -x
bash tools/gpu_ab.sh ah libqsb200_exp1.so libqsb200_exp2.so
QSB_LIB_PATH=qsearch_b200/libqsb200_exp1.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o /tmp/ah_n4 python tools/prof_case.py --what resid --qubits 4 --instances 20000 --reps 2 > gpurun_out/ah_ncu_full.log 2>&1
ncu -i /tmp/ah_n4.ncu-rep --page raw --csv > gpurun_out/ah_raw_n4.csv 2>/dev/null
ncu -i /tmp/ah_n4.ncu-rep --page source --csv > gpurun_out/ah_src_n4.csv 2>/dev/null
