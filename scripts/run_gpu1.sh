#!/bin/bash
# GPU round-trip 1: batch parity tests, then the whole GPU suite, then memcheck on one small batch case.
mkdir -p gpurun_out
python -m pytest tests/test_batch_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/batch_tests.log 2>&1
echo "batch exit $?" >> gpurun_out/batch_tests.log
tail -30 gpurun_out/batch_tests.log
python -m pytest tests -m gpu -q --timeout 900 --deselect tests/test_batch_gpu.py > gpurun_out/all_tests.log 2>&1
echo "all exit $?" >> gpurun_out/all_tests.log
tail -5 gpurun_out/all_tests.log
timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_batch_gpu.py -m gpu -q -x -k "projection_functions or exact_column" > gpurun_out/batch_memcheck.log 2>&1
tail -5 gpurun_out/batch_memcheck.log
