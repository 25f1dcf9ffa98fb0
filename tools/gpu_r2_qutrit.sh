#!/bin/bash
# qutrit path on the B200: tests, the two-qutrit synthesis cases of examples/qutrit_csum.py, kernel throughput
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -x -q -k "qutrit" 2>&1 | tail -25 | tee gpurun_out/t_pytest_qutrit.log
timeout 600 python tools/synth_bench.py --cases qt_csum,qt_bswap --reps 10 --seed0 3000 --timeout 100 > gpurun_out/t_synth_qutrit_10seeds.log 2>&1
cut -c1-60,215-330 gpurun_out/t_synth_qutrit_10seeds.log | tail -21
timeout 300 python tools/qutrit_bench.py 2>&1 | tee gpurun_out/t_qutrit_bench.log
