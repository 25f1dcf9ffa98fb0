#!/bin/bash
# validation of the CTA-per-instance hot kernel (uniform control flow, no WARPSYNC.COLLECTIVE) + projector shortcut at n=5:
# smoke, full GPU suite, benches, reference arm, launch list, ncu of the hot kernel n=4 / n=5, LM profile, QFT5 synthesis
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/an_smoke.log 2>&1; tail -4 gpurun_out/an_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/an_pytest_gpu.log 2>&1; tail -9 gpurun_out/an_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/an_bench_n4.json 2> gpurun_out/an_bench_n4.err; cut -c1-200 gpurun_out/an_bench_n4.json; tail -3 gpurun_out/an_bench_n4.err
timeout 600 python bench.py --steps 3 --warmup 3 --qubits 5 --vectors 200 --e2e-instances 4096 > gpurun_out/an_bench_n5.json 2> gpurun_out/an_bench_n5.err; cut -c1-200 gpurun_out/an_bench_n5.json; tail -3 gpurun_out/an_bench_n5.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/an_bench_reference.json 2>&1; cut -c1-200 gpurun_out/an_bench_reference.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/an_launches_bench_n4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/an_ncu_list.log 2>&1
for n in 4 5; do
  inst=$([ $n = 4 ] && echo 20000 || echo 4000)
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:qsb_k_warp_jac -c 1 -o /tmp/an_n$n python tools/prof_case.py --what resid --qubits $n --instances $inst --reps 2 > gpurun_out/an_ncu_full_n$n.log 2>&1
  ncu -i /tmp/an_n$n.ncu-rep --page raw --csv > gpurun_out/an_raw_n$n.csv 2>/dev/null
  ncu -i /tmp/an_n$n.ncu-rep --page source --csv > gpurun_out/an_src_n$n.csv 2>/dev/null
done
QSB_LM_PROFILE=1 timeout 200 python tools/solve_bench.py --method lm > gpurun_out/an_solve_lm2_prof.log 2>&1; grep -E "instances=[0-9]{2,}|case" gpurun_out/an_solve_lm2_prof.log | cut -c1-250
timeout 500 python tools/synth_bench.py --cases qft3,toffoli,qft4_leap --reps 2 > gpurun_out/an_synth_defaults.log 2>&1; cut -c1-250 gpurun_out/an_synth_defaults.log
timeout 700 python tools/synth_bench.py --cases qft5_yorktown_leap --reps 2 --timeout 300 > gpurun_out/an_synth_qft5.log 2>&1; cut -c1-300 gpurun_out/an_synth_qft5.log
