#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 400 $TR --nproc-per-node 2 --master-port 29531 tests/mgpu_worker.py > gpurun_out/mgpu2_worker.log 2>&1
echo "worker2 exit $?" | tee -a gpurun_out/mgpu2_worker.log
grep -v "^\*\*\*\|OMP_NUM" gpurun_out/mgpu2_worker.log | tail -12
timeout 400 $TR --nproc-per-node 2 --master-port 29532 scripts/bench_configs45.py > gpurun_out/configs45_n2.jsonl 2> gpurun_out/configs45_n2.err
echo "configs45 exit $?"; grep '^{' gpurun_out/configs45_n2.jsonl | cut -c1-500; tail -3 gpurun_out/configs45_n2.err
