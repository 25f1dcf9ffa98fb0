#!/bin/bash
# two-GPU call with the current kernels: torchrun bench n=4 and n=5 (weak scaling, one process per GPU), 2-device context test
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two_devices" > gpurun_out/aa_pytest2.log 2>&1; tail -3 gpurun_out/aa_pytest2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/aa_bench_n4_2gpu.json 2> gpurun_out/aa_bench_2gpu.err; cut -c1-220 gpurun_out/aa_bench_n4_2gpu.json; tail -3 gpurun_out/aa_bench_2gpu.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 --no-cpu > gpurun_out/aa_bench_n5_2gpu.json 2> gpurun_out/aa_bench_n5_2gpu.err; cut -c1-220 gpurun_out/aa_bench_n5_2gpu.json
