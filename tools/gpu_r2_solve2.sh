#!/bin/bash
# body-frame LM after a kernel change: GPU suite, phase profile, timings with and without the straggler rule: tools/gpu_r2_solve2.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; tail -4 gpurun_out/${tag}_pytest_gpu.log
QSB_LM_PROFILE=1 timeout 300 python tools/solve_bench.py --method lm > gpurun_out/${tag}_solve_lm3_prof.log 2>&1
grep profile gpurun_out/${tag}_solve_lm3_prof.log | cut -c1-300
for c in 0 2; do
  timeout 300 python tools/solve_bench.py --method lm --reps 3 --straggler $c > gpurun_out/${tag}_solve_lm3_strag$c.log 2>&1
  grep case gpurun_out/${tag}_solve_lm3_strag$c.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['case'], d['wall_ms_all'], d['status_counts'], 'iters max/sum', d['iters_max'], d['iters_sum'], 'us/iter', d['us_per_iter_latency'], 'best %.4g' % d['best_f'])
"
done
