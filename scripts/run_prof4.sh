#!/bin/bash
mkdir -p gpurun_out
python scripts/bench_config2.py | tee gpurun_out/config2.json
python scripts/prof_case.py --mode log --n 16384 --m 16384 --iters 3 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:log_pp_sweep -s 4 -c 1 -f -o gpurun_out/prof_log16k_pp python scripts/prof_case.py --mode log --n 16384 --m 16384 --iters 3 > gpurun_out/ncu2.log 2>&1; echo "ncu log exit $?"
SIZES=16384 ncu --set full --clock-control none --import-source on -k regex:emp_tc_lse -s 4 -c 1 -f -o gpurun_out/prof_emp_tc_16k_v3 python scripts/bench_emp.py > gpurun_out/ncu_emp.log 2>&1; echo "ncu emp exit $?"
