#!/bin/bash
mkdir -p gpurun_out
python scripts/diag_api_overhead.py 2>&1 | grep "config 1" | tee gpurun_out/r1s.log
python -m pytest tests/test_dense_gpu.py tests/test_solvers_gpu.py -x -q -m gpu 2>&1 | tail -2 | tee -a gpurun_out/r1s.log
