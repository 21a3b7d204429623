#!/bin/bash
# round 1, session o: batched solver with the register-resident fast path: parity tests + config 5 timing (on / off)
mkdir -p gpurun_out
python -m pytest tests/test_batch_gpu.py -x -q -m gpu > gpurun_out/r1o_batch_tests.log 2>&1; echo "batch tests rc=$?" | tee -a gpurun_out/r1o.log
tail -3 gpurun_out/r1o_batch_tests.log
echo "== regtile on" | tee -a gpurun_out/r1o.log
python scripts/bench_batch.py 2>&1 | tail -1 | tee -a gpurun_out/r1o.log
echo "== regtile off" | tee -a gpurun_out/r1o.log
B2OT_BATCH_REGTILE=0 python scripts/bench_batch.py 2>&1 | tail -1 | tee -a gpurun_out/r1o.log
