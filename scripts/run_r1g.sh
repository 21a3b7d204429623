#!/bin/bash
# round 1, session g: full-size property tests, epilogue + unbalanced benches, ncu of the epilogue kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fullsize_gpu.py -x -q -m gpu > gpurun_out/r1g_fullsize_tests.log 2>&1; echo "fullsize tests rc=$?" | tee -a gpurun_out/r1g_summary.log
tail -15 gpurun_out/r1g_fullsize_tests.log
python -m pytest tests/test_solvers_gpu.py -x -q -m gpu > gpurun_out/r1g_solvers_tests.log 2>&1; echo "solvers tests rc=$?" | tee -a gpurun_out/r1g_summary.log
tail -3 gpurun_out/r1g_solvers_tests.log
python scripts/bench_value.py 16384 > gpurun_out/r1g_bench_value.jsonl 2>&1; echo "bench_value rc=$?" | tee -a gpurun_out/r1g_summary.log
cat gpurun_out/r1g_bench_value.jsonl
python scripts/bench_unbalanced.py 65536 50 > gpurun_out/r1g_bench_unbalanced.json 2>&1; echo "bench_unbalanced rc=$?" | tee -a gpurun_out/r1g_summary.log
cat gpurun_out/r1g_bench_unbalanced.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:plan_value_kernel -s 6 -c 2 -f -o gpurun_out/r1g_plan_value python scripts/bench_value.py 16384 > gpurun_out/r1g_ncu.log 2>&1; echo "ncu rc=$?" | tee -a gpurun_out/r1g_summary.log
