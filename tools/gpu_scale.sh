#!/bin/bash
# weak-scaling bench lines on N GPUs of one box (one process per GPU, torchrun): tools/gpu_scale.sh <tag> <N>
tag=$1; N=$2
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_gpus.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu > gpurun_out/${tag}_bench_n4_${N}gpu.json 2> gpurun_out/${tag}_bench_n4_${N}gpu.err; cut -c1-220 gpurun_out/${tag}_bench_n4_${N}gpu.json; tail -3 gpurun_out/${tag}_bench_n4_${N}gpu.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus $N --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 --no-cpu > gpurun_out/${tag}_bench_n5_${N}gpu.json 2> gpurun_out/${tag}_bench_n5_${N}gpu.err; cut -c1-220 gpurun_out/${tag}_bench_n5_${N}gpu.json; tail -3 gpurun_out/${tag}_bench_n5_${N}gpu.err
