#!/bin/bash
mkdir -p gpurun_out
for lib in libqsb200.so libqsb200_lb4.so libqsb200_lb5.so; do
for q in 4 5; do
QSB_LIB_PATH=$PWD/qsearch_b200/$lib timeout 120 python tools/prof_case.py --what resid --qubits $q --instances 40000 --reps 4 2>&1 | sed "s/^/$lib /" | tee -a gpurun_out/p_occupancy_ab.log
done; done
