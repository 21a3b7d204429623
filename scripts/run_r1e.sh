#!/bin/bash
# round 1, session e: new solvers tests, full GPU suite, epilogue micro-benchmark
mkdir -p gpurun_out
python -m pytest tests/test_solvers_gpu.py -x -q -m gpu > gpurun_out/r1e_solvers_tests.log 2>&1; echo "solvers tests rc=$?" | tee -a gpurun_out/r1e_summary.log
tail -15 gpurun_out/r1e_solvers_tests.log
python scripts/bench_value.py 16384 > gpurun_out/r1e_bench_value.jsonl 2>&1; echo "bench_value rc=$?" | tee -a gpurun_out/r1e_summary.log
cat gpurun_out/r1e_bench_value.jsonl
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/r1e_all_tests.log 2>&1; echo "all tests rc=$?" | tee -a gpurun_out/r1e_summary.log
tail -5 gpurun_out/r1e_all_tests.log
