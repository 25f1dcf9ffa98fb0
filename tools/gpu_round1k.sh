#!/bin/bash
# QFT5 Yorktown LEAP: beams x starts trade-off at a fixed 240 s budget
mkdir -p gpurun_out
for cfg in "1 6" "2 3" "4 2"; do set -- $cfg
timeout 330 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 1 --starts $1 --beams $2 --timeout 240 2>&1 | tail -1 | tee -a gpurun_out/k_qft5_beams.log
done
timeout 200 python tools/synth_bench.py --cases qft4_leap --reps 2 --starts 1 --beams 8 --timeout 120 2>&1 | tail -2 | tee -a gpurun_out/k_qft5_beams.log
timeout 200 python tools/synth_bench.py --cases qft4_leap --reps 2 --starts 4 --beams 4 --timeout 120 2>&1 | tail -2 | tee -a gpurun_out/k_qft5_beams.log
