#!/bin/bash
# config 5 (QFT5 on IBM Yorktown, LEAP) with one frontier sharded over 1 / 2 / 4 / 8 GPUs of one box: tools/gpu_r2_multi.sh <tag> [reps]
tag=$1; reps=${2:-4}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_gpus.txt 2>&1
ngpu=$(nvidia-smi -L | wc -l)
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two_devices" > gpurun_out/${tag}_pytest_two_devices.log 2>&1; tail -2 gpurun_out/${tag}_pytest_two_devices.log
run() { name=$1; devs=$2; shift 2; QSB200_DEVICES=$devs timeout 900 python tools/synth_bench.py --cases qft5_yorktown_leap --reps $reps --timeout 120 "$@" > gpurun_out/${tag}_synth_qft5_${name}.log 2>&1
  grep case gpurun_out/${tag}_synth_qft5_${name}.log | python -c "
import sys, json
rows = [json.loads(l) for l in sys.stdin]
print('$name', 'devices', rows[0]['devices'] if rows else '?', 'wall', [r['wall_s'] for r in rows], 'cnots', [r['cnots'] for r in rows], 'exp', [r['stats']['expansions'] for r in rows], 'conv', [r['distance'] < 1e-10 for r in rows])"; }
run d1 0
[ $ngpu -ge 2 ] && run d2 0,1
[ $ngpu -ge 4 ] && run d4 0,1,2,3
if [ $ngpu -ge 8 ]; then
  run d8 0,1,2,3,4,5,6,7
  run d8_b16 0,1,2,3,4,5,6,7 --beams 16 --fstarts 8
  run d8_b24 0,1,2,3,4,5,6,7 --beams 24 --fstarts 6
  QSB200_DEVICES=0,1,2,3,4,5,6,7 timeout 300 python tools/synth_bench.py --cases qft4_leap --reps 3 --min-depth 4 > gpurun_out/${tag}_synth_qft4_d8.log 2>&1; grep case gpurun_out/${tag}_synth_qft4_d8.log | cut -c1-260
fi
