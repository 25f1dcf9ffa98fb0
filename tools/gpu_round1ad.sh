#!/bin/bash
# A/B of the projector-type derivative shortcut in the hot kernel: base = library built before the change
set -x
mkdir -p gpurun_out
for n in 4 5; do
  inst=$([ $n = 4 ] && echo 100000 || echo 20000)
  for lib in base new; do
    p=$([ $lib = base ] && echo qsearch_b200/libqsb200_base.so || echo qsearch_b200/libqsb200.so)
    QSB_LIB_PATH=$p timeout 200 python tools/prof_case.py --what resid --qubits $n --instances $inst --reps 6 2>&1 | sed "s/^/[$lib n=$n] /"
    QSB_LIB_PATH=$p timeout 200 python tools/prof_case.py --what matjac --qubits $n --instances $inst --reps 4 2>&1 | sed "s/^/[$lib n=$n] /"
  done
done > gpurun_out/ad_ab_projector_shortcut.log 2>&1
cat gpurun_out/ad_ab_projector_shortcut.log
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/ad_pytest_parity.log 2>&1; tail -5 gpurun_out/ad_pytest_parity.log
