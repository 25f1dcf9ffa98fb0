#!/bin/bash
# stop rules in qsb_k_lm2 (circuits beyond qsb_k_lm3's shared-memory factor): full GPU suite, QFT5/Yorktown synthesis, 8 seeds: tools/gpu_r2_strag3.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/${tag}_pytest_gpu.log
QSB_TRACE_SOLVES=1 timeout 900 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 8 --timeout 120 > gpurun_out/${tag}_synth_qft5.log 2>&1
python - $tag <<'PY'
import json,sys,re
walls=[]
for line in open("gpurun_out/%s_synth_qft5.log" % sys.argv[1]):
    if line.startswith("{"):
        d=json.loads(line); print(d["case"], d["rep"], d["wall_s"], d["cnots"], "%.1e" % d["distance"], d["stats"])
    m=re.search(r"wall_ms=([0-9.]+)", line)
    if m: walls.append(float(m.group(1)))
walls.sort(reverse=True); print("slowest solves (ms):", walls[:8], "of", len(walls))
PY
