#!/bin/bash
# A/B of hot-kernel builds through QSB_LIB_PATH: tools/gpu_ab.sh <tag> <lib> [<lib> ...]   (libs relative to qsearch_b200/)
tag=$1; shift
mkdir -p gpurun_out
for rep in 1 2; do
for n in ${QSB_AB_N:-4 5}; do
  inst=$([ $n = 4 ] && echo 100000 || echo 20000)
  for lib in "$@"; do
    QSB_LIB_PATH=qsearch_b200/$lib timeout 200 python tools/prof_case.py --what resid --qubits $n --instances $inst --circuits 100 --reps 6 2>&1 | sed "s/^/[$lib n=$n resid] /"
    QSB_LIB_PATH=qsearch_b200/$lib timeout 200 python tools/prof_case.py --what matjac --qubits $n --instances $inst --circuits 100 --reps 4 2>&1 | sed "s/^/[$lib n=$n matjac] /"
  done
done
done > gpurun_out/${tag}_ab.log 2>&1
grep "evals/s" gpurun_out/${tag}_ab.log
