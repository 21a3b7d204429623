#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gromov_gpu.py -q -m gpu 2>&1 | grep -E "AssertionError|Error|^FAILED|passed|failed" | head -20 | tee gpurun_out/r2a.log
timeout 300 python tests/perf_gromov.py 1000 10 200 2>&1 | tail -1 | tee -a gpurun_out/r2a.log
timeout 300 python tests/perf_gromov.py 4000 5 200 2>&1 | tail -1 | tee -a gpurun_out/r2a.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stab_rebuild_kernel -s 5 -c 1 -f -o gpurun_out/r2a_stab_rebuild python scripts/bench_stabilized.py 16384 30 > gpurun_out/r2a_ncu.log 2>&1; echo "ncu rc=$?"
