#!/bin/bash
# straggler-cap activation fix: repeatability of bench.py's frontier_rules and the QFT5 / QFT4 synthesis tails: tools/gpu_r2_strag2.sh <tag>
tag=$1
mkdir -p gpurun_out
bash tools/gpu_r2_mu2.sh ${tag}
timeout 600 python -m pytest tests -m gpu -x -q -k "stop_rules or synthesis or leap or cnot_counts or qutrit" 2>&1 | tail -3 | tee gpurun_out/${tag}_pytest_gpu.log
timeout 200 python tools/synth_bench.py --cases qft4_leap --reps 5 --min-depth 4 --timeout 100 > gpurun_out/${tag}_synth_qft4.log 2>&1
timeout 900 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 8 --timeout 120 > gpurun_out/${tag}_synth_qft5.log 2>&1
python - $tag <<'PY'
import json,sys
for f in ("synth_qft4","synth_qft5"):
    for line in open("gpurun_out/%s_%s.log" % (sys.argv[1], f)):
        if line.startswith("{"):
            d=json.loads(line); print(d["case"], d["rep"], d["wall_s"], d["cnots"], "%.1e" % d["distance"], d["stats"])
PY
