#!/bin/bash
mkdir -p gpurun_out
for lib in libqsb200_v1.so libqsb200.so libqsb200_v1.so libqsb200.so; do
  QSB_LIB_PATH=qsearch_b200/$lib timeout 300 python tools/e2e_case.py --reps 6 2>&1 | sed "s/^/[$lib] /"
done > gpurun_out/as_e2e_ab.log 2>&1
cat gpurun_out/as_e2e_ab.log
