#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_stabilized_gpu.py -x -q -m gpu 2>&1 | tail -15 | tee gpurun_out/r1x.log
timeout 900 python -m pytest tests/test_dense_gpu.py tests/test_solvers_gpu.py -x -q -m gpu 2>&1 | tail -2 | tee -a gpurun_out/r1x.log
timeout 300 python scripts/bench_stabilized.py 16384 500 2>&1 | tail -4 | tee -a gpurun_out/r1x.log
