#!/bin/bash
# resident solver of small dense problems: dense test suite + bench config 1 (with and without it)
mkdir -p gpurun_out/r2q
timeout 900 python -m pytest tests/test_dense_gpu.py -m gpu -x -q > gpurun_out/r2q/pytest_dense.log 2>&1; echo "pytest dense rc=$?"; tail -3 gpurun_out/r2q/pytest_dense.log
timeout 600 python bench.py --config 1 > gpurun_out/r2q/bench_c1.json 2> gpurun_out/r2q/bench_c1.err; echo "bench c1 rc=$?"; cut -c1-250 gpurun_out/r2q/bench_c1.json; tail -3 gpurun_out/r2q/bench_c1.err
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2q/pytest_all.log 2>&1; echo "pytest all rc=$?"; tail -3 gpurun_out/r2q/pytest_all.log
