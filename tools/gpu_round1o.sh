#!/bin/bash
# LEAP + re-optimising post-processor end to end (configs[2] and configs[4] as the reference's examples run them)
mkdir -p gpurun_out
timeout 600 python tools/synth_bench.py --cases qft4_leap --reps 3 --reopt 7 --timeout 200 2>&1 | tail -3 | tee gpurun_out/o_synth_qft4_reopt.log
timeout 900 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 1 --reopt 7 --timeout 400 2>&1 | tail -1 | tee gpurun_out/o_synth_qft5_reopt.log
