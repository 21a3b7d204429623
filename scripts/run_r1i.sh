#!/bin/bash
# round 1, session i (2 GPUs): multi-GPU parity worker incl. solve_sharded, headline bench at N=2 (peer exchange)
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_worker.py > gpurun_out/r1i_mgpu_worker.log 2>&1
echo "worker exit $?" | tee -a gpurun_out/r1i_mgpu_worker.log
tail -12 gpurun_out/r1i_mgpu_worker.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/r1i_bench_n2.json 2> gpurun_out/r1i_bench_n2.err
echo "bench exit $?"; tail -c 2500 gpurun_out/r1i_bench_n2.json; tail -3 gpurun_out/r1i_bench_n2.err
