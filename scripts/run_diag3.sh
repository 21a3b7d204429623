#!/bin/bash
mkdir -p gpurun_out
for g in fused1 cluster perf; do
  timeout 420 python tests/gpu_diag.py --group $g --quick > gpurun_out/diag_$g.log 2>&1
  echo "exit $?" >> gpurun_out/diag_$g.log
  grep -E "BAD|PERF|bad cases|exit|rror" gpurun_out/diag_$g.log
done
for mode in log knopp; do
ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum --clock-control none -k regex:fused_sweep -s 4 -c 1 python scripts/prof_case.py --mode $mode --n 16384 --m 16384 --iters 3 2>&1 | grep -E "inst_executed|duration"
done
