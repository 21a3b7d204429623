#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gromov_batch_gpu.py -q -m gpu 2>&1 | grep -E "Error|^FAILED|passed|failed|^E  " | head -30 | tee gpurun_out/r2b.log
