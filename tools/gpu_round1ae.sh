#!/bin/bash
# ncu of the hot kernel with the projector-type shortcut (slower than the base in the A/B): where does the time go
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o /tmp/ae_n4 python tools/prof_case.py --what resid --qubits 4 --instances 20000 --reps 2 > gpurun_out/ae_ncu_full.log 2>&1
ncu -i /tmp/ae_n4.ncu-rep --page raw --csv > gpurun_out/ae_raw_n4.csv 2>/dev/null
ncu -i /tmp/ae_n4.ncu-rep --page source --csv > gpurun_out/ae_src_n4.csv 2>/dev/null
python tools/ncu_digest.py gpurun_out/ae_raw_n4.csv gpurun_out/ae_src_n4.csv > gpurun_out/ae_ncu_digest_n4.txt 2>&1
head -50 gpurun_out/ae_ncu_digest_n4.txt
