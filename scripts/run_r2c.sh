#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_empirical_gpu.py -m gpu -q 2>&1 | tail -3 | tee gpurun_out/r2c2.log
for r in 1 2 3; do timeout 600 python -m pytest tests/test_empirical_gpu.py -m gpu -q -k "tc_vs_oracle or tc_golden" 2>&1 | tail -1 | tee -a gpurun_out/r2c2.log; done
SIZES=16384,32768,65536 timeout 300 python scripts/bench_emp.py 2>&1 | grep -v simt | cut -c1-110 | tee -a gpurun_out/r2c2.log
