#!/bin/bash
# e2e A/B on one box: bench.py with the previous library build (v1: 4-warp CTAs) and the current one
set -x
mkdir -p gpurun_out
cp qsearch_b200/libqsb200.so /tmp/cur.so
for lib in v1 cur v1 cur; do
  if [ $lib = v1 ]; then cp qsearch_b200/libqsb200_v1.so qsearch_b200/libqsb200.so; else cp /tmp/cur.so qsearch_b200/libqsb200.so; fi
  timeout 600 python bench.py --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib', 'value %.4e'%d['value'], 'e2e %.4e'%d['e2e']['value'], 'hs_e2e %.4e'%d['hs_form']['e2e_value'])"
done > gpurun_out/ar_e2e_ab.log 2>&1
cp /tmp/cur.so qsearch_b200/libqsb200.so
cat gpurun_out/ar_e2e_ab.log
