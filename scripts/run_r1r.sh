#!/bin/bash
# round 1, session r: CUDA-graph replay of iteration chunks: parity (whole GPU suite), config 1 latency, headline sanity
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r1r_all_tests.log 2>&1; echo "all gpu tests rc=$?" | tee gpurun_out/r1r.log
tail -3 gpurun_out/r1r_all_tests.log
python scripts/diag_api_overhead.py 2>&1 | grep "its=\|config 1" | tee -a gpurun_out/r1r.log
echo "== graphs off" | tee -a gpurun_out/r1r.log
B2OT_GRAPH=0 python scripts/diag_api_overhead.py 2>&1 | grep "its=\|config 1" | tee -a gpurun_out/r1r.log
python bench.py --steps 100 --warmup 10 > gpurun_out/r1r_bench.json 2> gpurun_out/r1r_bench.err; echo "bench rc=$?" | tee -a gpurun_out/r1r.log
cat gpurun_out/r1r_bench.json | cut -c1-1600
