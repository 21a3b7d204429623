#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/ab_finalize.jsonl
for r in 1 2 3; do
  python ab_prev/scripts/bench_unbalanced.py 65536 100 balanced_only 2>&1 | tail -1 | sed 's/^/{"lib": "prev (6751f3a)", "run": /; s/$/}/' >> gpurun_out/ab_finalize.jsonl
  python scripts/bench_unbalanced.py 65536 100 balanced_only 2>&1 | tail -1 | sed 's/^/{"lib": "new (stabilized finalize)", "run": /; s/$/}/' >> gpurun_out/ab_finalize.jsonl
done
cat gpurun_out/ab_finalize.jsonl
nvidia-smi --query-gpu=clocks.sm,clocks_throttle_reasons.sw_power_cap,power.draw,temperature.gpu --format=csv
