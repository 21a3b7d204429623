#!/bin/bash
# round 2, first GPU call: full GPU test-suite (incl. the 2-ranks-on-one-GPU product test) + every config of bench.py
mkdir -p gpurun_out/r2a
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2a/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a/pytest.log
tail -5 gpurun_out/r2a/pytest.log
for c in 3 2 1 5 4; do
  timeout 600 python bench.py --config $c > gpurun_out/r2a/bench_c$c.json 2> gpurun_out/r2a/bench_c$c.err; echo "bench c$c rc=$?"
  tail -c 600 gpurun_out/r2a/bench_c$c.json
done
timeout 300 python bench.py --impl reference --config 3 --steps 20 --warmup 3 > gpurun_out/r2a/ref_c3.json 2>&1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
