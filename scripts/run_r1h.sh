#!/bin/bash
# round 1, session h: A/B of the headline sweep with / without the unbalanced row-exponent branch, unbalanced bench
# with the MUFU row exponent, epilogue bench
mkdir -p gpurun_out
python -m pytest tests/test_solvers_gpu.py -x -q -m gpu -k "unbalanced" > gpurun_out/r1h_unb_tests.log 2>&1; echo "unbalanced tests rc=$?" | tee -a gpurun_out/r1h_summary.log
tail -3 gpurun_out/r1h_unb_tests.log
for rep in 1 2 3; do
  python scripts/bench_unbalanced.py 65536 100 >> gpurun_out/r1h_ab_pow.jsonl 2>&1
done
echo "--- pow"; cat gpurun_out/r1h_ab_pow.jsonl
python scripts/bench_value.py 16384 > gpurun_out/r1h_bench_value.jsonl 2>&1
cat gpurun_out/r1h_bench_value.jsonl
