#!/bin/bash
# ncu --set full of the online-cost sweep, variants v1 (single pass, clamp), v2 (no clamp), v5 (row operand in TMEM)
mkdir -p gpurun_out/r2j
for v in v1 v2 v5; do
  export B2OT_LIB=pot_b200/libb2ot_$v.so
  SIZES=32768 VARIANTS=tc timeout 600 ncu --set full --clock-control none --import-source on -k regex:emp_tc_lse -s 6 -c 1 -f \
    -o gpurun_out/r2j/emp_$v python scripts/bench_emp.py > gpurun_out/r2j/ncu_$v.log 2>&1
  echo "$v ncu rc=$?"; tail -2 gpurun_out/r2j/ncu_$v.log
done
ls -la gpurun_out/r2j
