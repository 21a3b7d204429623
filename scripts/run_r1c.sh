#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_dense_gpu.py -m gpu -x -q > gpurun_out/pytest_dense.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_dense.log
for pp in 1 0; do echo "== LOGPP $pp"
B2OT_LOGPP=$pp python scripts/prof_case.py --mode log --n 16384 --m 16384 --iters 100 2>&1 | grep PERF
B2OT_LOGPP=$pp python scripts/prof_case.py --mode log --n 4096 --m 4096 --iters 200 2>&1 | grep PERF
B2OT_LOGPP=$pp python scripts/prof_case.py --mode log --n 65536 --m 65536 --iters 20 2>&1 | grep PERF
done
