#!/bin/bash
# round 2, second GPU call: full GPU test-suite after centring of the tensor-core operands, gradients, batch helpers
mkdir -p gpurun_out/r2b
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2b/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b/pytest.log
tail -15 gpurun_out/r2b/pytest.log
timeout 300 python bench.py --config 4 --n 65536 --no-cpu --no-e2e > gpurun_out/r2b/bench_c4_65536.json 2> gpurun_out/r2b/bench_c4_65536.err
tail -c 400 gpurun_out/r2b/bench_c4_65536.json
