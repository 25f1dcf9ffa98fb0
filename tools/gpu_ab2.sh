#!/bin/bash
# A/B on the bench's own shape (100 circuits): tools/gpu_ab2.sh <tag> <lib> [<lib> ...]
tag=$1; shift
mkdir -p gpurun_out
for n in 4 5; do
  inst=$([ $n = 4 ] && echo 100000 || echo 20000)
  for lib in "$@"; do
    QSB_LIB_PATH=qsearch_b200/$lib timeout 200 python tools/prof_case.py --what resid --qubits $n --instances $inst --circuits 100 --reps 6 2>&1 | sed "s/^/[$lib n=$n] /"
  done
done > gpurun_out/${tag}_ab.log 2>&1
cat gpurun_out/${tag}_ab.log
