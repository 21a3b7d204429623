#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_empirical_gpu.py -m gpu -x -q > gpurun_out/pytest_emp.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/pytest_emp.log
SIZES=${SIZES:-16384,65536} timeout 300 python scripts/bench_emp.py 2>&1 | grep -v simt | tee gpurun_out/bench_emp.jsonl | cut -c1-330
