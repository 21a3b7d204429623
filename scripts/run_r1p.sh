#!/bin/bash
mkdir -p gpurun_out
B=296 timeout 600 ncu --set full --clock-control none --import-source on -k regex:batch_solve_kernel -s 11 -c 1 -f -o gpurun_out/r1p_batch_log python scripts/bench_batch.py > gpurun_out/r1p_ncu.log 2>&1; echo "ncu rc=$?"
B=296 timeout 600 ncu --set full --clock-control none --import-source on -k regex:batch_solve_kernel -s 4 -c 1 -f -o gpurun_out/r1p_batch_knopp python scripts/bench_batch.py > gpurun_out/r1p_ncu2.log 2>&1; echo "ncu rc=$?"
