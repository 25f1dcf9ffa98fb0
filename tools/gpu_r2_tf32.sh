#!/bin/bash
# TF32 Gram option of the body-frame LM: tools/gpu_r2_tf32.sh <tag>
tag=$1
mkdir -p gpurun_out
for tf in 0 1; do
  QSB_LM_PROFILE=1 timeout 300 python tools/solve_bench.py --method lm --tf32 $tf > gpurun_out/${tag}_solve_tf$tf_prof.log 2>&1
  grep profile gpurun_out/${tag}_solve_tf$tf_prof.log | awk 'NR%2==0' | cut -c1-260
  timeout 300 python tools/solve_bench.py --method lm --reps 3 --tf32 $tf > gpurun_out/${tag}_solve_tf$tf.log 2>&1
  grep case gpurun_out/${tag}_solve_tf$tf.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print('tf32=$tf', d['case'], d['wall_ms_all'], d['status_counts'], 'iters max/sum', d['iters_max'], d['iters_sum'], 'us/iter', d['us_per_iter_latency'], 'best %.4g' % d['best_f'])
"
done
