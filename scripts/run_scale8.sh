#!/bin/bash
# 8-GPU round-trip: parity worker on 8 ranks, then the headline bench at N = 8 and 4 with both exchanges.
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 tests/mgpu_worker.py > gpurun_out/mgpu8_worker.log 2>&1
echo "worker8 exit $?" | tee -a gpurun_out/mgpu8_worker.log
grep -v "^\*\*\*\|OMP_NUM" gpurun_out/mgpu8_worker.log | tail -12
for n in 8 4; do
for ex in peer nccl; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $n --steps 200 --warmup 20 --exchange $ex --no-e2e > gpurun_out/bench_n${n}_$ex.json 2> gpurun_out/bench_n${n}_$ex.err
  echo "bench n=$n $ex exit $?"; grep '^{' gpurun_out/bench_n${n}_$ex.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['value'], d['ms_per_step'], d['roofline']['launch_ms'], d['roofline']['frac'])"; tail -2 gpurun_out/bench_n${n}_$ex.err
done
done
