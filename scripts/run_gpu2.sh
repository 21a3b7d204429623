#!/bin/bash
# GPU round-trip 2: empirical parity tests + batch tests, then timing of the batch config-5 shape.
mkdir -p gpurun_out
python -m pytest tests/test_empirical_gpu.py -m gpu -q --timeout 600 > gpurun_out/emp_tests.log 2>&1
echo "emp exit $?" >> gpurun_out/emp_tests.log
tail -40 gpurun_out/emp_tests.log
python -m pytest tests/test_batch_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/batch_tests.log 2>&1
echo "batch exit $?" >> gpurun_out/batch_tests.log
tail -3 gpurun_out/batch_tests.log
python scripts/bench_batch.py > gpurun_out/bench_batch.log 2>&1
cat gpurun_out/bench_batch.log
