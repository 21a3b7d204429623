#!/bin/bash
mkdir -p gpurun_out
for cs in ${CSLIST:-2 4 1}; do
export B2OT_EMP_CLUSTER=$cs
echo "== cluster $cs"
if [ -z "$NOTEST" ]; then timeout 600 python -m pytest tests/test_empirical_gpu.py -m gpu -x -q > gpurun_out/pytest_emp_$cs.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/pytest_emp_$cs.log; fi
SIZES=${SIZES:-16384,65536} timeout 300 python scripts/bench_emp.py 2>&1 | grep -v simt | tee gpurun_out/bench_emp_$cs.jsonl | cut -c1-250
done
