#!/bin/bash
# 4-warp / two-CTAs-per-SM shape of the body-frame LM kernel against the 8-warp shape, same box: tools/gpu_r2_w4.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/${tag}_pytest_gpu.log
for w in 8 4 8 4; do
  for tf in 0 1; do
    echo "== QSB_LM_WARPS=$w tf32=$tf"
    QSB_LM_WARPS=$w timeout 200 python tools/solve_bench.py --method lm --only "n4 d10 4x250" --reps 3 --tf32 $tf 2>&1 | grep case | cut -c1-330
  done
done > gpurun_out/${tag}_solve_w4_ab.log 2>&1
cat gpurun_out/${tag}_solve_w4_ab.log
timeout 300 python bench.py --no-cpu --steps 3 > gpurun_out/${tag}_bench_n4.json 2> gpurun_out/${tag}_bench_n4.err
python - $tag <<'PY'
import json,sys
d=json.load(open("gpurun_out/%s_bench_n4.json" % sys.argv[1])); sf=d["solve_form"]
for k in ("reference_semantics","frontier_rules"): print(k, round(sf[k]["solves_per_s"]), sf[k]["wall_s"], sf[k]["us_per_lm_iteration_latency"], sf[k]["best_distance"])
PY
