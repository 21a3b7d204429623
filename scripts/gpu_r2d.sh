#!/bin/bash
# 2 GPUs: multi-rank parity worker (new put kernel, un-gated flags, slot-reuse stress) + config 3 / 2 / 4 / 5 bench lines at N = 2
mkdir -p gpurun_out/r2d
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/r2d/pytest_mgpu.log 2>&1; echo "rc=$?" >> gpurun_out/r2d/pytest_mgpu.log
tail -8 gpurun_out/r2d/pytest_mgpu.log
for c in 3 2 4 5; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --config $c --no-cpu > gpurun_out/r2d/bench_c${c}_n2.json 2> gpurun_out/r2d/bench_c${c}_n2.err
  echo "bench c$c n2 rc=$?"; tail -c 300 gpurun_out/r2d/bench_c${c}_n2.json
done
