#!/bin/bash
# usage: scripts/run_diag.sh  (on the GPU box)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.sm,clocks.max.sm --format=csv > gpurun_out/diag_gpu.txt 2>&1
for g in split fused1 cluster perf; do
  echo "=== group $g ===" 
  timeout 420 python tests/gpu_diag.py --group $g > gpurun_out/diag_$g.log 2>&1
  echo "exit $?" >> gpurun_out/diag_$g.log
  tail -n 60 gpurun_out/diag_$g.log
done
