#!/bin/bash
# ncu --set full of the LM solver kernel (n = 4, 1000 instances, reference semantics), raw + source pages kept as CSV: tools/gpu_r2_ncu_lm3.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 100 python tools/solve_bench.py --method lm --only "n4 d10 4x250" > gpurun_out/${tag}_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qsb_k_lm3 -s 1 -c 1 -o gpurun_out/${tag}_ncu_lm3 python tools/solve_bench.py --method lm --only "n4 d10 4x250" > gpurun_out/${tag}_ncu_lm3.log 2>&1
ncu -i gpurun_out/${tag}_ncu_lm3.ncu-rep --page raw --csv > gpurun_out/${tag}_raw_lm3_n4_d10_4x250.csv 2>/dev/null
ncu -i gpurun_out/${tag}_ncu_lm3.ncu-rep --page source --csv > gpurun_out/${tag}_src_lm3_n4_d10_4x250.csv 2>/dev/null
python tools/ncu_digest.py raw gpurun_out/${tag}_raw_lm3_n4_d10_4x250.csv > gpurun_out/${tag}_ncu_full_lm3_n4_d10_4x250.txt 2>&1
python tools/ncu_digest.py source gpurun_out/${tag}_src_lm3_n4_d10_4x250.csv >> gpurun_out/${tag}_ncu_full_lm3_n4_d10_4x250.txt 2>&1
ls -la gpurun_out/${tag}_*
rm -f gpurun_out/${tag}_ncu_lm3.ncu-rep
head -40 gpurun_out/${tag}_ncu_full_lm3_n4_d10_4x250.txt | cut -c1-170
