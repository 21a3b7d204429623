#!/bin/bash
mkdir -p gpurun_out
for g in fused1 cluster perf; do
  timeout 420 python tests/gpu_diag.py --group $g > gpurun_out/diag_$g.log 2>&1
  echo "exit $?" >> gpurun_out/diag_$g.log
  grep -E "BAD|PERF|bad cases|exit|rror" gpurun_out/diag_$g.log
done
