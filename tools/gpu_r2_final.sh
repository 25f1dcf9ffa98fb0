#!/bin/bash
# round-2 evidence on one GPU: suite, bench lines, reference arm, latency, D2H, launch list, ncu of the solver kernel: tools/gpu_r2_final.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${tag}_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/${tag}_bench_n4.json 2> gpurun_out/${tag}_bench_n4.err; cut -c1-160 gpurun_out/${tag}_bench_n4.json; tail -2 gpurun_out/${tag}_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/${tag}_bench_n5.json 2> gpurun_out/${tag}_bench_n5.err; cut -c1-160 gpurun_out/${tag}_bench_n5.json; tail -2 gpurun_out/${tag}_bench_n5.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2>&1; cut -c1-160 gpurun_out/${tag}_bench_reference.json
timeout 300 python tools/latency_bench.py > gpurun_out/${tag}_latency.log 2>&1; cat gpurun_out/${tag}_latency.log | cut -c1-200
nvcc -O2 -o /tmp/d2h_bench tools/d2h_bench.cu > /dev/null 2>&1 && /tmp/d2h_bench 0 512 8 > gpurun_out/${tag}_d2h.log 2>&1; cat gpurun_out/${tag}_d2h.log
timeout 300 python tools/solve_bench.py --method lm --reps 5 > gpurun_out/${tag}_solve_lm3.log 2>&1; grep case gpurun_out/${tag}_solve_lm3.log | cut -c1-200
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench_n4.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-solve > gpurun_out/${tag}_ncu_list.log 2>&1
for c in "n4 d10 4x250" "n5 d12 6x4"; do
  f=$(echo $c | tr ' ' '_')
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_lm3 -s 1 -c 1 -o gpurun_out/${tag}_ncu_lm3_$f python tools/solve_bench.py --method lm --only "$c" > gpurun_out/${tag}_ncu_lm3_$f.log 2>&1
  ncu -i gpurun_out/${tag}_ncu_lm3_$f.ncu-rep --page raw --csv > gpurun_out/${tag}_raw_lm3_$f.csv 2>/dev/null
  ncu -i gpurun_out/${tag}_ncu_lm3_$f.ncu-rep --page source --csv > gpurun_out/${tag}_src_lm3_$f.csv 2>/dev/null
  python tools/ncu_digest.py raw gpurun_out/${tag}_raw_lm3_$f.csv > gpurun_out/${tag}_ncu_full_lm3_$f.txt 2>&1
  python tools/ncu_digest.py source gpurun_out/${tag}_src_lm3_$f.csv >> gpurun_out/${tag}_ncu_full_lm3_$f.txt 2>&1
  rm -f gpurun_out/${tag}_ncu_lm3_$f.ncu-rep gpurun_out/${tag}_src_lm3_$f.csv
  head -30 gpurun_out/${tag}_ncu_full_lm3_$f.txt | cut -c1-160
done
