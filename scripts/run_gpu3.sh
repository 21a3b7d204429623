#!/bin/bash
# GPU round-trip 3: tensor-core empirical sweep bring-up (guarded by timeouts), then timing at growing sizes.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_empirical_gpu.py -m gpu -q -x --timeout 120 -k "tc" > gpurun_out/emp_tc_tests.log 2>&1
echo "tc tests exit $?" >> gpurun_out/emp_tc_tests.log
tail -40 gpurun_out/emp_tc_tests.log
timeout 600 python scripts/bench_emp.py > gpurun_out/bench_emp.log 2>&1
echo "bench_emp exit $?"; tail -20 gpurun_out/bench_emp.log
