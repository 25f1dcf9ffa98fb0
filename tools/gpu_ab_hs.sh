#!/bin/bash
# A/B of HS-kernel builds through QSB_LIB_PATH: tools/gpu_ab_hs.sh <tag> <lib> [<lib> ...]
tag=$1; shift
mkdir -p gpurun_out
for n in 3 4 5; do
  inst=$([ $n = 5 ] && echo 100000 || echo 200000)
  for lib in "$@"; do
    QSB_LIB_PATH=qsearch_b200/$lib timeout 200 python tools/prof_case.py --what hs --qubits $n --instances $inst --reps 6 2>&1 | sed "s/^/[$lib n=$n] /"
  done
done > gpurun_out/${tag}_ab_hs.log 2>&1
cat gpurun_out/${tag}_ab_hs.log
