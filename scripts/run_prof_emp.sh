#!/bin/bash
mkdir -p gpurun_out
SIZES=16384 ncu --set full --clock-control none --import-source on -k regex:emp_tc_lse -s 4 -c 1 -f -o gpurun_out/prof_emp_tc_16k_v2 python scripts/bench_emp.py > gpurun_out/ncu_emp.log 2>&1; echo "ncu emp exit $?"
