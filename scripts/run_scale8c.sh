#!/bin/bash
# 8-GPU refresh at the end of the round: parity worker, headline bench at N = 8 (peer, with e2e), configs 4 and 5 at N = 8.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29521 tests/mgpu_worker.py > gpurun_out/mgpu8_worker.log 2>&1
echo "worker8 exit $?" | tee -a gpurun_out/mgpu8_worker.log
grep -v "^\*\*\*\|OMP_NUM" gpurun_out/mgpu8_worker.log | tail -8
timeout 400 $TR --nproc-per-node 8 --master-port 29522 bench.py --gpus 8 --steps 200 --warmup 20 --exchange peer > gpurun_out/bench_n8_peer.json 2> gpurun_out/bench_n8_peer.err
echo "bench n=8 peer exit $?"; grep '^{' gpurun_out/bench_n8_peer.json | cut -c1-160; tail -2 gpurun_out/bench_n8_peer.err
timeout 400 $TR --nproc-per-node 8 --master-port 29525 scripts/bench_configs45.py > gpurun_out/configs45_n8.jsonl 2> gpurun_out/configs45_n8.err
echo "configs45 exit $?"; grep '^{' gpurun_out/configs45_n8.jsonl | cut -c1-400; tail -3 gpurun_out/configs45_n8.err
