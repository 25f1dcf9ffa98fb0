#!/bin/bash
# TF32 Gram option after the template split: suite, solver timings, QFT5 synthesis with / without: tools/gpu_r2_tf32b.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${tag}_pytest_gpu.log
for tf in 0 1; do
  timeout 300 python tools/solve_bench.py --method lm --reps 3 --tf32 $tf > gpurun_out/${tag}_solve_tf$tf.log 2>&1
  grep case gpurun_out/${tag}_solve_tf$tf.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print('tf32=$tf', d['case'], d['wall_ms_all'], 'iters max/sum', d['iters_max'], d['iters_sum'], 'us/iter', d['us_per_iter_latency'], 'best %.4g' % d['best_f'])
"
done
for tf in 1 0; do
  timeout 700 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 5 --seed0 1000 --timeout 120 --tf32 $tf > gpurun_out/${tag}_synth_qft5_tf$tf.log 2>&1
  grep case gpurun_out/${tag}_synth_qft5_tf$tf.log | python -c "
import sys, json
rows = [json.loads(l) for l in sys.stdin]
print('qft5 tf32=$tf wall', [r['wall_s'] for r in rows], 'cnots', [r['cnots'] for r in rows], 'exp', [r['stats']['expansions'] for r in rows], 'conv', [r['distance'] < 1e-10 for r in rows])"
done
timeout 200 python tools/synth_bench.py --cases qft4_leap --reps 5 --seed0 1000 --min-depth 4 > gpurun_out/${tag}_synth_qft4.log 2>&1; grep case gpurun_out/${tag}_synth_qft4.log | cut -c100-230
