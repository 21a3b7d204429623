#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_empirical_gpu.py -m gpu -x -q > gpurun_out/pytest_emp_f16.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_emp_f16.log
for k in f16 tf32; do echo "== kind $k"; B2OT_EMP_KIND=$k SIZES=16384,65536 timeout 300 python scripts/bench_emp.py 2>&1 | grep -v simt | cut -c1-120; done
for d in 2 4; do echo "== f16 dbg $d"; B2OT_EMP_DBG=$d SIZES=32768 timeout 300 python scripts/bench_emp.py 2>&1 | grep -v simt | cut -c50-100; done
B2OT_EMP_DBG=0 python scripts/trace_emp.py | sed -n 1,3p\;20,32p
