#!/bin/bash
mkdir -p gpurun_out
export B2OT_OCC=2
timeout 300 python tests/gpu_diag.py --group cluster --quick 2>&1 | grep -E "BAD|bad cases|rror|config log 33"
timeout 300 python tests/gpu_diag.py --group fused1 --quick 2>&1 | grep -E "BAD|bad cases|rror"
for mode in knopp log; do
python scripts/prof_case.py --mode $mode --n 16384 --m 16384 --iters 100 --label "occ2" 2>&1 | grep -E "PERF|rror"
python scripts/prof_case.py --mode $mode --n 8192 --m 65536 --iters 50 --label "occ2" 2>&1 | grep -E "PERF|rror"
done
