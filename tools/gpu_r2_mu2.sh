#!/bin/bash
# bench.py solve_form (three seeded frontiers, n = 4) under LM damping-schedule variants: tools/gpu_r2_mu2.sh <tag>
tag=$1
mkdir -p gpurun_out
i=0
for v in "QSB_X=0" "QSB_X=1" "QSB_X=2"; do
  i=$((i+1))
  echo "== $v"
  env $v timeout 200 python bench.py --no-cpu --steps 2 > gpurun_out/${tag}_bench_v$i.json 2>/dev/null
  python tools/print_solve_form.py gpurun_out/${tag}_bench_v$i.json
done 2>&1 | tee gpurun_out/${tag}_mu_bench.log
rm -f gpurun_out/${tag}_bench_v*.json
