#!/bin/bash
# synthesis tuning on one GPU: tools/gpu_r2_synth.sh <tag>
tag=$1
mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python tools/synth_bench.py "$@" > gpurun_out/${tag}_synth_${name}.log 2>&1; grep case gpurun_out/${tag}_synth_${name}.log | cut -c1-330; tail -2 gpurun_out/${tag}_synth_${name}.log | grep -i error; }
run 3q --cases qft3,toffoli,fredkin,peres,or,qft3_ring,toffoli_ring --reps 3
run qft4 --cases qft4_leap --reps 3 --min-depth 4
run qft5_auto --cases qft5_yorktown_leap --reps 3 --timeout 300
run qft5_s1 --cases qft5_yorktown_leap --reps 3 --fstarts 1 --timeout 300
run qft5_s3 --cases qft5_yorktown_leap --reps 3 --fstarts 3 --timeout 300
run qft5_s3_nostrag --cases qft5_yorktown_leap --reps 2 --fstarts 3 --straggler 0 --timeout 300
