#!/bin/bash
# dense GPU tests + log/knopp timing after the packed-fp32x2 row pass
mkdir -p gpurun_out
python -m pytest tests/test_dense_gpu.py -m gpu -x -q > gpurun_out/pytest_dense.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_dense.log
for mode in log knopp; do python scripts/prof_case.py --mode $mode --n 16384 --m 16384 --iters 100; done 2>&1 | grep -v "^$" | tee gpurun_out/perf_log16k.log
python scripts/prof_case.py --mode log --n 65536 --m 65536 --iters 20 2>&1 | tee -a gpurun_out/perf_log16k.log
