#!/bin/bash
# synthesis end to end with the current solver: configs[2] and configs[4], with and without the re-optimising post-processor
mkdir -p gpurun_out
timeout 600 python tools/synth_bench.py --cases qft4_leap --reps 3 --reopt 7 --timeout 200 2>&1 | tail -3 | tee gpurun_out/z_synth_qft4_reopt.log
timeout 1500 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 3 --reopt 7 --timeout 400 2>&1 | tail -3 | tee gpurun_out/z_synth_qft5_reopt.log
