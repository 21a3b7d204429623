#!/bin/bash
# hybrid batch kernel: parity tests + config 5 bench (both domains)
mkdir -p gpurun_out/r2f
timeout 900 python -m pytest tests/test_batch_gpu.py tests/test_fullsize_gpu.py tests/test_batchops_gpu.py tests/test_gromov_batch_gpu.py -m gpu -q -k "not 65536" > gpurun_out/r2f/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2f/pytest.log
tail -12 gpurun_out/r2f/pytest.log
timeout 300 python bench.py --config 5 --no-cpu > gpurun_out/r2f/bench_c5.json 2> gpurun_out/r2f/bench_c5.err; tail -c 1500 gpurun_out/r2f/bench_c5.json
timeout 300 python scripts/bench_batch.py > gpurun_out/r2f/bench_batch.log 2>&1; tail -12 gpurun_out/r2f/bench_batch.log
