#!/bin/bash
# round 1, session n: ncu source-level capture of the batched on-chip solver (config 5 shape, small batch)
mkdir -p gpurun_out
python scripts/bench_batch.py > gpurun_out/r1n_batch.json 2>&1; cat gpurun_out/r1n_batch.json
B=296 timeout 600 ncu --set full --clock-control none --import-source on -k regex:batch_solve_kernel -s 4 -c 2 -f -o gpurun_out/r1n_batch python scripts/bench_batch.py > gpurun_out/r1n_ncu.log 2>&1; echo "ncu rc=$?"
