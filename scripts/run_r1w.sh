#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_dense_gpu.py tests/test_solvers_gpu.py tests/test_fullsize_gpu.py -x -q -m gpu 2>&1 | tail -2 | tee gpurun_out/r1w.log
python scripts/bench_value.py 16384 2>&1 | grep "plan_kernel\|cost_transform\|loss_kernel" | tee -a gpurun_out/r1w.log
python scripts/bench_value.py 49152 2>&1 | grep "float32" | grep "plan_kernel\|cost_transform\|loss_kernel\|plan+linear+kl" | tee -a gpurun_out/r1w.log
