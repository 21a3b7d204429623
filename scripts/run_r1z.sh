#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -6 | tee gpurun_out/r1z_tests.log
timeout 600 python bench.py > gpurun_out/r1z_bench.json 2> gpurun_out/r1z_bench.err; tail -c 2500 gpurun_out/r1z_bench.json
timeout 300 python scripts/diag_api_overhead.py 2>&1 | tail -12 | tee gpurun_out/r1z_api_overhead.log
timeout 300 python scripts/bench_stabilized.py 1000 1000 2>&1 | tail -2 | tee gpurun_out/r1z_stab_1000.log
