#!/bin/bash
# GPU suite, then the no-allocation test alone in a fresh process, then bench lines: tools/gpu_r2_check.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/${tag}_pytest_gpu.log 2>&1; tail -12 gpurun_out/${tag}_pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${tag}_pytest_noalloc_alone.log 2>&1; tail -3 gpurun_out/${tag}_pytest_noalloc_alone.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log
timeout 900 python bench.py > gpurun_out/${tag}_bench_n4.json 2> gpurun_out/${tag}_bench_n4.err; cut -c1-300 gpurun_out/${tag}_bench_n4.json; tail -3 gpurun_out/${tag}_bench_n4.err
