#!/bin/bash
# 2-GPU round-trip: multi-GPU parity worker, then the headline bench at N=2 with both exchanges.
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_worker.py > gpurun_out/mgpu_worker.log 2>&1
echo "worker exit $?" | tee -a gpurun_out/mgpu_worker.log
tail -25 gpurun_out/mgpu_worker.log
for ex in peer nccl; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 100 --warmup 10 --exchange $ex --no-e2e > gpurun_out/bench_n2_$ex.json 2> gpurun_out/bench_n2_$ex.err
  echo "bench $ex exit $?"; tail -c 1500 gpurun_out/bench_n2_$ex.json; tail -3 gpurun_out/bench_n2_$ex.err
done
