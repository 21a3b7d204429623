#!/bin/bash
# round 1, session f: unbalanced solver tests, headline bench sanity after the sweep-kernel change
mkdir -p gpurun_out
python -m pytest tests/test_solvers_gpu.py -x -q -m gpu -k "unbalanced" > gpurun_out/r1f_unb_tests.log 2>&1; echo "unbalanced tests rc=$?" | tee -a gpurun_out/r1f_summary.log
tail -30 gpurun_out/r1f_unb_tests.log
python bench.py --steps 50 --warmup 5 > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err; echo "bench rc=$?" | tee -a gpurun_out/r1f_summary.log
cat gpurun_out/r1f_bench.json
