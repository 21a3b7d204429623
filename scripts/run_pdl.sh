#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_dense_gpu.py tests/test_batch_gpu.py -m gpu -x -q > gpurun_out/pytest_dense.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_dense.log
for pdl in 1 0; do echo "== PDL $pdl"; B2OT_PDL=$pdl python bench.py --no-e2e --no-cpu --steps 100 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['launch_ms'], d['roofline']['frac'])"; done
for pdl in 1 0; do echo "== PDL $pdl n=16384"; B2OT_PDL=$pdl python scripts/prof_case.py --mode knopp --n 16384 --m 16384 --iters 200 2>&1 | grep PERF; B2OT_PDL=$pdl python scripts/prof_case.py --mode knopp --n 8192 --m 65536 --iters 200 2>&1 | grep PERF;  done
