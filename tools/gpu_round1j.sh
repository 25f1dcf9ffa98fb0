#!/bin/bash
# two-GPU call: torchrun bench (weak scaling, one process per GPU), reference arm under torchrun, 2-device context test
set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two_devices" > gpurun_out/j_pytest2.log 2>&1; tail -3 gpurun_out/j_pytest2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/j_bench_n4_2gpu.json 2> gpurun_out/j_bench_2gpu.err; cat gpurun_out/j_bench_n4_2gpu.json; tail -5 gpurun_out/j_bench_2gpu.err
timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 > gpurun_out/j_bench_n4_1gpu.json 2> gpurun_out/j_bench_1gpu.err; cat gpurun_out/j_bench_n4_1gpu.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/j_bench_ref_2gpu.json 2>&1; tail -2 gpurun_out/j_bench_ref_2gpu.json
