#!/bin/bash
# Final checks of the round-2 build (mode-dependent LM damping schedule): tools/gpu_r2_final3.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/${tag}_pytest_gpu.log
timeout 200 python tools/solve_bench.py --method lm --reps 3 > gpurun_out/${tag}_solve_lm3.log 2>&1
timeout 200 python tools/solve_bench.py --method lm --reps 3 --straggler 2 --sibling 1e-10 --tf32 1 > gpurun_out/${tag}_solve_lm3_frontier.log 2>&1
grep -h case gpurun_out/${tag}_solve_lm3.log gpurun_out/${tag}_solve_lm3_frontier.log | cut -c1-140
timeout 400 python bench.py --steps 5 > gpurun_out/${tag}_bench_n4.json 2> gpurun_out/${tag}_bench_n4.err
timeout 400 python bench.py --steps 5 --qubits 5 > gpurun_out/${tag}_bench_n5.json 2> gpurun_out/${tag}_bench_n5.err
python tools/print_solve_form.py gpurun_out/${tag}_bench_n4.json; python tools/print_solve_form.py gpurun_out/${tag}_bench_n5.json
timeout 300 python tools/synth_bench.py --cases qft3,toffoli,fredkin_ring,qft4_leap --reps 5 --min-depth 4 --timeout 100 > gpurun_out/${tag}_synth.log 2>&1
timeout 600 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 5 --timeout 120 > gpurun_out/${tag}_synth_qft5.log 2>&1
python - $tag <<'PY'
import json,sys
for f in ("synth","synth_qft5"):
    for line in open("gpurun_out/%s_%s.log" % (sys.argv[1], f)):
        if line.startswith("{"):
            d=json.loads(line); print(d["case"], d["rep"], d["wall_s"], d["cnots"], "%.1e" % d["distance"], d["stats"])
PY
