#!/bin/bash
mkdir -p gpurun_out
for g in fused1 cluster; do
  timeout 420 python tests/gpu_diag.py --group $g --quick > gpurun_out/diag_$g.log 2>&1
  echo "exit $?" >> gpurun_out/diag_$g.log
  grep -E "BAD|bad cases|rror" gpurun_out/diag_$g.log
done
echo "--- swapped warp counts (knopp 16 / log 8) via env ---"
B2OT_WARPS=16 timeout 300 python tests/gpu_diag.py --group cluster --quick 2>&1 | grep -E "BAD|bad cases|rror"
B2OT_WARPS=8 timeout 300 python tests/gpu_diag.py --group fused1 --quick 2>&1 | grep -E "BAD|bad cases|rror"
for w in 8 16; do
for mode in knopp log; do
B2OT_WARPS=$w python scripts/prof_case.py --mode $mode --n 16384 --m 16384 --iters 100 --label "warps=$w" 2>&1 | grep PERF
B2OT_WARPS=$w python scripts/prof_case.py --mode $mode --n 8192 --m 65536 --iters 50 --label "warps=$w" 2>&1 | grep PERF
done
done
