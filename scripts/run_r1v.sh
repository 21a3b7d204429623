#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_batch_gpu.py -x -q -m gpu > gpurun_out/r1v_batch_tests.log 2>&1; echo "batch tests rc=$?"
tail -15 gpurun_out/r1v_batch_tests.log
