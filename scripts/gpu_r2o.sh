#!/bin/bash
# checkpoint of the cleaned-up online-cost sweep: full GPU test suite, bench lines of config 4 (65536 and 262144 on one GPU),
# ncu --set full of one half-sweep at 32768^2
mkdir -p gpurun_out/r2o
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2o/pytest_all.log 2>&1; echo "pytest all rc=$?"; tail -3 gpurun_out/r2o/pytest_all.log
timeout 600 python bench.py --config 4 --n 65536 --no-cpu > gpurun_out/r2o/bench_c4_n65536.json 2> gpurun_out/r2o/bench_c4_n65536.err; echo "bench 65536 rc=$?"; cut -c1-300 gpurun_out/r2o/bench_c4_n65536.json
timeout 900 python bench.py --config 4 > gpurun_out/r2o/bench_c4.json 2> gpurun_out/r2o/bench_c4.err; echo "bench c4 rc=$?"; cut -c1-300 gpurun_out/r2o/bench_c4.json
SIZES=32768 VARIANTS=tc timeout 600 ncu --set full --clock-control none --import-source on -k regex:emp_tc_lse -s 6 -c 1 -f \
  -o gpurun_out/r2o/emp_final python scripts/bench_emp.py > gpurun_out/r2o/ncu_final.log 2>&1; echo "ncu rc=$?"
