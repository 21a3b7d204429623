#!/bin/bash
# hybrid batch kernel v2: A/B of 4 vs 2 rows per group in the log domain, parity tests on the default
mkdir -p gpurun_out/r2h
timeout 300 python scripts/bench_batch.py > gpurun_out/r2h/bench_rg4.log 2>&1; tail -1 gpurun_out/r2h/bench_rg4.log | cut -c1-700
B2OT_LIB=pot_b200/libb2ot_rg2.so timeout 300 python scripts/bench_batch.py > gpurun_out/r2h/bench_rg2.log 2>&1; tail -1 gpurun_out/r2h/bench_rg2.log | cut -c1-700
timeout 900 python -m pytest tests/test_batch_gpu.py tests/test_fullsize_gpu.py tests/test_gromov_batch_gpu.py -m gpu -q -k "not 65536" > gpurun_out/r2h/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2h/pytest.log
tail -4 gpurun_out/r2h/pytest.log
B2OT_LIB=pot_b200/libb2ot_rg2.so timeout 900 python -m pytest tests/test_batch_gpu.py -m gpu -q > gpurun_out/r2h/pytest_rg2.log 2>&1; tail -2 gpurun_out/r2h/pytest_rg2.log
