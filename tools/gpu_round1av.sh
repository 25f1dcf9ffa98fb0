#!/bin/bash
# end-to-end synthesis with the round's final kernels: README 3-qubit set, QFT4 LEAP + reopt, QFT5 Yorktown LEAP (4 reps), stateprep
mkdir -p gpurun_out
timeout 300 python tools/synth_bench.py --cases qft3,toffoli,fredkin,peres,or,qft3_ring,toffoli_ring --reps 3 > gpurun_out/av_synth_3q.log 2>&1; cut -c1-220 gpurun_out/av_synth_3q.log | tail -22
timeout 300 python tools/synth_bench.py --cases qft4_leap --reps 3 --reopt 7 --timeout 120 > gpurun_out/av_synth_qft4_leap_reopt.log 2>&1; cut -c1-300 gpurun_out/av_synth_qft4_leap_reopt.log | tail -4
timeout 900 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 4 --timeout 200 > gpurun_out/av_synth_qft5_yorktown.log 2>&1; cut -c1-300 gpurun_out/av_synth_qft5_yorktown.log | tail -5
