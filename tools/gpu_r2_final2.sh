#!/bin/bash
# final state of round 2 on one GPU: smoke, suite, no-alloc test alone, bench lines, reference arm, ncu of the TF32 solver kernel
tag=$1
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; tail -3 gpurun_out/${tag}_smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${tag}_pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${tag}_pytest_noalloc_alone.log 2>&1; tail -1 gpurun_out/${tag}_pytest_noalloc_alone.log
timeout 600 python bench.py > gpurun_out/${tag}_bench_n4.json 2> gpurun_out/${tag}_bench_n4.err; cut -c1-160 gpurun_out/${tag}_bench_n4.json; tail -2 gpurun_out/${tag}_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/${tag}_bench_n5.json 2> gpurun_out/${tag}_bench_n5.err; cut -c1-160 gpurun_out/${tag}_bench_n5.json; tail -2 gpurun_out/${tag}_bench_n5.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2>&1; cut -c1-160 gpurun_out/${tag}_bench_reference.json
c="n5 d30 6x24"; f=$(echo $c | tr ' ' '_')
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qsb_k_lm3 -s 1 -c 1 -o gpurun_out/${tag}_ncu_lm3tf32_$f python tools/solve_bench.py --method lm --tf32 1 --only "$c" > gpurun_out/${tag}_ncu_lm3tf32_$f.log 2>&1
ncu -i gpurun_out/${tag}_ncu_lm3tf32_$f.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
ncu -i gpurun_out/${tag}_ncu_lm3tf32_$f.ncu-rep --page source --csv > gpurun_out/${tag}_src.csv 2>/dev/null
python tools/ncu_digest.py raw gpurun_out/${tag}_raw.csv > gpurun_out/${tag}_ncu_full_lm3tf32_$f.txt 2>&1
python tools/ncu_digest.py source gpurun_out/${tag}_src.csv >> gpurun_out/${tag}_ncu_full_lm3tf32_$f.txt 2>&1
rm -f gpurun_out/${tag}_ncu_lm3tf32_$f.ncu-rep gpurun_out/${tag}_src.csv gpurun_out/${tag}_raw.csv
head -32 gpurun_out/${tag}_ncu_full_lm3tf32_$f.txt | cut -c1-150
